// oracle/cvshim/list_arena.h -- TEST INFRASTRUCTURE ONLY.
//
// Pre-included (g++ -include) ahead of the reference's src/ORBextractor.cc when building
// oracle/_ref.  Purpose: canonicalise the one allocator-dependent step of the reference.
//
// ORBextractor::DistributeOctTree sorts vector<pair<int, ExtractorNode*>>
// (src/ORBextractor.cc:682): ties in the key count are broken by the HEAP ADDRESS of the
// std::list node, so with glibc malloc the survivors depend on heap history (SURVEY.md 8 E3,
// experiment X6).  The canonical rule used by this project is "creation order" -- what any
// monotonic allocator gives.  We obtain it without editing the reference file: every
// `list<...>` token in the reference TU is re-pointed (macro below) at std::list with a
// thread-local bump allocator whose addresses strictly increase with allocation order and
// are never reused until orb_list_arena_reset().
#ifndef ORACLE_LIST_ARENA_H
#define ORACLE_LIST_ARENA_H

// Everything the reference TU includes is pulled in first, so the macro at the bottom only
// ever rewrites tokens of the two reference files themselves.
#include <algorithm>
#include <iostream>
#include <iterator>
#include <list>
#include <vector>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <sys/mman.h>

#include "cvshim.h"

namespace orb_arena {

struct Arena {
    char* base = nullptr;
    size_t used = 0;
    size_t cap = 0;
    bool enabled = true;  // false -> plain malloc (reproduces the allocator-dependent behaviour)
};
inline Arena& tls() {
    static thread_local Arena a;
    return a;
}
inline void reset() {
    Arena& a = tls();
    // The pages stay mapped between frames (a per-frame madvise would put page faults and mmap-lock traffic
    // into the timed CPU arm that a stock build with malloc does not pay); only an unusually large frame's
    // worth is given back.
    if (a.base && a.used > ((size_t)256 << 20)) madvise(a.base, a.used, MADV_DONTNEED);
    a.used = 0;
}
inline void* bump(size_t n) {
    Arena& a = tls();
    if (!a.base) {
        a.cap = (size_t)1 << 30;  // virtual reservation, touched lazily
        void* p = mmap(nullptr, a.cap, PROT_READ | PROT_WRITE,
                       MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
        if (p == MAP_FAILED) { perror("orb_arena mmap"); abort(); }
        a.base = (char*)p;
    }
    n = (n + 15) & ~(size_t)15;
    if (a.used + n > a.cap) { fprintf(stderr, "orb_arena: exhausted\n"); abort(); }
    void* r = a.base + a.used;
    a.used += n;
    return r;
}

template <typename T>
struct BumpAlloc {
    typedef T value_type;
    BumpAlloc() {}
    template <typename U>
    BumpAlloc(const BumpAlloc<U>&) {}
    T* allocate(size_t n) {
        if (!tls().enabled) return (T*)malloc(n * sizeof(T));
        return (T*)bump(n * sizeof(T));
    }
    void deallocate(T* p, size_t) {
        Arena& a = tls();
        if ((char*)p >= a.base && (char*)p < a.base + a.cap && a.base) return;  // never reused
        free(p);
    }
    template <typename U>
    bool operator==(const BumpAlloc<U>&) const { return true; }
    template <typename U>
    bool operator!=(const BumpAlloc<U>&) const { return false; }
};

}  // namespace orb_arena

namespace std {
template <typename T>
using orb_arena_list = std::list<T, orb_arena::BumpAlloc<T> >;
}

#define list orb_arena_list

#endif  // ORACLE_LIST_ARENA_H
