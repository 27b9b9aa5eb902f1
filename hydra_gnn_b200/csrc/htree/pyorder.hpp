// Iteration-order model of the two CPython containers the reference's H-tree construction leans on.
//
// The reference builds H-trees with networkx (neural_tree/construct.py, generate_junction_tree_hierarchies.py).
// Node numbering of the result depends on the iteration order of Python sets (tie-breaks of the maximum
// cardinality search, node order of sub-graph views, `{... for ...}` comprehensions) and of insertion-ordered
// dicts (networkx adjacency).  To be bit-exact the native builder reproduces both:
//   * PySet  -- CPython 3.12 Objects/setobject.c: open addressing, 9 linear probes then the
//               i*5+1+perturb recurrence, dummies on discard (the LAST dummy met is reused), resize at
//               fill*5 >= mask*3 to used*4, clean re-insertion in old table order, set_merge's three paths;
//   * tuple_hash -- CPython >= 3.8 Objects/tupleobject.c (xxHash-style lanes); hash(int) == int for the
//               non-negative node ids used here.
// Validated against the running interpreter by tests/test_htree.py (random op sequences through hg_pyset_trace).
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

namespace hg_htree {

inline int64_t tuple_hash(const int* items, size_t n) {
    const uint64_t P1 = 11400714785074694791ULL, P2 = 14029467366897019727ULL, P5 = 2870177450012600261ULL;
    uint64_t acc = P5;
    for (size_t i = 0; i < n; ++i) {
        acc += (uint64_t)(int64_t)items[i] * P2;  // lane = hash(int) = int (ids are >= 0)
        acc = (acc << 31) | (acc >> 33);
        acc *= P1;
    }
    acc += (uint64_t)n ^ (P5 ^ 3527539ULL);
    if (acc == (uint64_t)-1) return 1546275796;
    return (int64_t)acc;
}

class PySet {
public:
    static constexpr int64_t kUnused = -1, kDummy = -2;  // keys are >= 0
    struct Entry { int64_t hash; int64_t key; };

    PySet() : t_(8, Entry{0, kUnused}) {}

    size_t size() const { return used_; }
    bool empty() const { return used_ == 0; }

    void add(int64_t key, int64_t hash) {
        const size_t mask = mask_;
        uint64_t perturb = (uint64_t)hash;
        size_t i = (size_t)perturb & mask;
        long freeslot = -1;
        for (;;) {
            int probes = (i + kLinearProbes <= mask) ? kLinearProbes : 0;
            size_t j = i;
            for (;;) {
                Entry& e = t_[j];
                if (e.key == kUnused) {
                    if (freeslot < 0) {
                        e = Entry{hash, key};
                        ++fill_; ++used_;
                        if (fill_ * 5 >= mask * 3) resize(used_ > 50000 ? used_ * 2 : used_ * 4);
                    } else {
                        t_[freeslot] = Entry{hash, key};
                        ++used_;
                    }
                    return;
                }
                if (e.key == kDummy) freeslot = (long)j;
                else if (e.hash == hash && e.key == key) return;
                ++j;
                if (probes-- == 0) break;
            }
            perturb >>= 5;
            i = (i * 5 + 1 + perturb) & mask;
        }
    }
    void add(int key) { add(key, key); }

    bool has(int64_t key, int64_t hash) const { return find(key, hash) >= 0; }
    bool has(int key) const { return find(key, key) >= 0; }

    void discard(int64_t key, int64_t hash) {
        long j = find(key, hash);
        if (j >= 0) { t_[j] = Entry{-1, kDummy}; --used_; }
    }
    void discard(int key) { discard(key, key); }

    // set.update(other_set) / set(other_set): Objects/setobject.c set_merge
    void merge(const PySet& o) {
        if (&o == this || o.used_ == 0) return;
        if ((fill_ + o.used_) * 5 >= mask_ * 3) resize((used_ + o.used_) * 2);
        if (fill_ == 0 && mask_ == o.mask_ && o.fill_ == o.used_) {
            t_ = o.t_; fill_ = o.fill_; used_ = o.used_;
            return;
        }
        if (fill_ == 0) {
            fill_ = used_ = o.used_;
            for (const Entry& e : o.t_) if (e.key >= 0) insert_clean(t_, mask_, e);
            return;
        }
        for (size_t i = 0; i <= o.mask_; ++i) if (o.t_[i].key >= 0) add(o.t_[i].key, o.t_[i].hash);
    }

    // a & b: iterates the smaller operand (ties: the right one), keeps keys found in the other
    static PySet intersection(const PySet& a, const PySet& b) {
        const PySet *so = &a, *other = &b;
        if (other->used_ > so->used_) { const PySet* tmp = so; so = other; other = tmp; }
        PySet r;
        for (const Entry& e : other->t_) if (e.key >= 0 && so->has(e.key, e.hash)) r.add(e.key, e.hash);
        return r;
    }

    template <class F> void each(F&& f) const {  // table order == Python iteration order
        for (size_t i = 0; i <= mask_; ++i) if (t_[i].key >= 0) f(t_[i].key);
    }
    std::vector<int> items() const {
        std::vector<int> v; v.reserve(used_);
        each([&](int64_t k) { v.push_back((int)k); });
        return v;
    }

private:
    static constexpr int kLinearProbes = 9;
    std::vector<Entry> t_;
    size_t mask_ = 7, fill_ = 0, used_ = 0;

    long find(int64_t key, int64_t hash) const {
        const size_t mask = mask_;
        uint64_t perturb = (uint64_t)hash;
        size_t i = (size_t)perturb & mask;
        for (;;) {
            int probes = (i + kLinearProbes <= mask) ? kLinearProbes : 0;
            size_t j = i;
            for (;;) {
                const Entry& e = t_[j];
                if (e.key == kUnused) return -1;
                if (e.key >= 0 && e.hash == hash && e.key == key) return (long)j;
                ++j;
                if (probes-- == 0) break;
            }
            perturb >>= 5;
            i = (i * 5 + 1 + perturb) & mask;
        }
    }
    static void insert_clean(std::vector<Entry>& t, size_t mask, const Entry& e) {
        uint64_t perturb = (uint64_t)e.hash;
        size_t i = (size_t)perturb & mask;
        for (;;) {
            if (t[i].key == kUnused) { t[i] = e; return; }
            if (i + kLinearProbes <= mask)
                for (int j = 1; j <= kLinearProbes; ++j)
                    if (t[i + j].key == kUnused) { t[i + j] = e; return; }
            perturb >>= 5;
            i = (i * 5 + 1 + perturb) & mask;
        }
    }
    void resize(size_t minused) {
        size_t newsize = 8;
        while (newsize <= minused) newsize <<= 1;
        std::vector<Entry> nt(newsize, Entry{0, kUnused});
        for (const Entry& e : t_) if (e.key >= 0) insert_clean(nt, newsize - 1, e);
        t_.swap(nt);
        mask_ = newsize - 1;
        fill_ = used_;
    }
};

}  // namespace hg_htree
