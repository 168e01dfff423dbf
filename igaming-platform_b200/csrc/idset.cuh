// idset.cuh — device-resident set of 64-bit identifiers (open addressing, linear probing, 0 = empty
// slot).  Stands in for the Redis SETs blacklist:devices / :fingerprints / :ips
// (features/redis_store.go:245-247); membership = SISMEMBER (:272-278).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

struct igs_idset_view {
    const uint64_t* slots;   // capacity = mask + 1 (power of two), nullptr = empty set
    uint64_t mask;
};

__host__ __device__ __forceinline__ uint64_t igs_idset_mix(uint64_t x) {   // splitmix64 finaliser
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull;
    x ^= x >> 27; x *= 0x94d049bb133111ebull;
    x ^= x >> 31;
    return x;
}

// id 0 is the empty string: never a member (the reference skips empty identifiers, :271-279)
__device__ __forceinline__ bool igs_idset_has(const igs_idset_view& s, uint64_t id) {
    if (id == 0 || s.slots == nullptr) return false;
    uint64_t h = igs_idset_mix(id) & s.mask;
    for (;;) {
        const uint64_t v = __ldg(s.slots + h);
        if (v == id) return true;
        if (v == 0) return false;          // the table is never more than half full
        h = (h + 1) & s.mask;
    }
}

// CheckBlacklist for one identifier triple (:267-293), in two phases so that the caller can put other
// work between them: begin() issues the first probe of each of the three sets (three independent random
// reads), end() inspects them and walks on only where the first slot held another id — at load factor
// <= 1/2 most lookups end on the first slot.
struct igs_idset_probe3 {
    uint64_t id[3], h[3], v[3];
    bool live[3];
};
__device__ __forceinline__ void igs_idset_probe_begin(const igs_idset_view s[3], uint64_t dev, uint64_t fp,
                                                      uint64_t ip, igs_idset_probe3* p) {
    p->id[0] = dev; p->id[1] = fp; p->id[2] = ip;          // order of IGS_BL_*
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        p->live[k] = p->id[k] != 0 && s[k].slots != nullptr;
        p->h[k] = igs_idset_mix(p->id[k]) & s[k].mask;
        p->v[k] = p->live[k] ? __ldg(s[k].slots + p->h[k]) : 0;
    }
}
__device__ __forceinline__ bool igs_idset_probe_end(const igs_idset_view s[3], const igs_idset_probe3& p) {
    bool hit = false;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        if (!p.live[k]) continue;
        uint64_t x = p.v[k], hh = p.h[k];
        while (x != p.id[k] && x != 0) {
            hh = (hh + 1) & s[k].mask;
            x = __ldg(s[k].slots + hh);
        }
        hit |= x == p.id[k];
    }
    return hit;
}

struct igs_engine;
// views of the engine's three blacklist sets, indexed by IGS_BL_*
void igs_blacklist_views(const igs_engine* e, igs_idset_view out[3]);
