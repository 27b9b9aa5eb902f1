"""Relation / node-type names of the reference's graph formats (dict keys and parameter names).

Values are the reference's (mp3d_dataset.py:21-26; neural_tree/construct.py:14-38); the
spellings ``"or_ro_rr"`` / ``"rr_ro_or"`` are the reference's own and must stay verbatim.
"""
EDGE_TYPES = [
    ("objects", "objects_to_objects", "objects"),
    ("rooms", "rooms_to_rooms", "rooms"),
    ("objects", "objects_to_rooms", "rooms"),
    ("rooms", "rooms_to_objects", "objects"),
]

HTREE_NODE_TYPES = ["object", "room", "object-room", "room-room"]
HTREE_EDGE_TYPES = [
    ("object", "o_to_or", "object-room"),
    ("object-room", "or_to_o", "object"),
    ("room", "r_to_or", "object-room"),
    ("object-room", "or_to_r", "room"),
    ("room", "r_to_rr", "room-room"),
    ("room-room", "rr_to_r", "room"),
    ("object-room", "or_ro_rr", "room-room"),
    ("room-room", "rr_ro_or", "object-room"),
    ("object-room", "or_to_or", "object-room"),
    ("room-room", "rr_to_rr", "room-room"),
]
HTREE_VIRTUAL_NODE_TYPES = ["object_virtual", "room_virtual"]
HTREE_INIT_EDGE_TYPES = [
    ("object_virtual", "ov_to_or", "object-room"),
    ("room_virtual", "rv_to_or", "object-room"),
    ("room_virtual", "rv_to_rr", "room-room"),
]
HTREE_POOL_EDGE_TYPES = [
    ("object", "o_to_ov", "object_virtual"),
    ("room", "r_to_rv", "room_virtual"),
]
