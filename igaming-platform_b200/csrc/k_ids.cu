// k_ids.cu — the blacklist sets on the device.
//
// Replaces the Redis SETs blacklist:devices / blacklist:fingerprints / blacklist:ips
// (features/redis_store.go:245-247): AddToBlacklist = SADD (:251-264), CheckBlacklist = three SISMEMBERs
// (:267-293).  A set is an open-addressing table of 64-bit ids (igs_hash_id of the string), kept at most
// half full; inserts are one atomicCAS per probe, lookups read-only (idset.cuh).  HBM-resident integer
// work: a probe is one random 8-byte read, ~1.5 probes per lookup at load factor 1/2.
#include <algorithm>

#include "idset.cuh"
#include "igs_internal.cuh"

namespace {

constexpr int kThreads = 256;

// inserts ids[0..n) (skipping 0); *added counts the ids that were not in the set yet
__global__ void k_idset_insert(unsigned long long* __restrict__ slots, uint64_t mask,
                               const uint64_t* __restrict__ ids, uint64_t n, unsigned long long* __restrict__ added) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint64_t id = ids[i];
    if (id == 0) return;
    uint64_t h = igs_idset_mix(id) & mask;
    for (;;) {
        const unsigned long long prev = atomicCAS(slots + h, 0ull, (unsigned long long)id);
        if (prev == 0ull) { atomicAdd(added, 1ull); return; }
        if (prev == id) return;
        h = (h + 1) & mask;
    }
}

__global__ void k_blacklist_check(igs_idset_view dev, igs_idset_view fp, igs_idset_view ip,
                                  const igs_tx_ids* __restrict__ ids, uint64_t n, uint8_t* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const igs_tx_ids t = ids[i];
    const igs_idset_view s[3] = {dev, fp, ip};
    igs_idset_probe3 p;
    igs_idset_probe_begin(s, t.device, t.fingerprint, t.ip, &p);
    out[i] = igs_idset_probe_end(s, p);
}

uint64_t table_cap_for(uint64_t count) {      // power of two, >= 2 * count, >= 1024
    uint64_t cap = 1024;
    while (cap < 2 * count) cap <<= 1;
    return cap;
}

}  // namespace

void igs_idset_free(igs_idset* s) {
    if (s->slots) cudaFree(s->slots);
    *s = igs_idset();
}

// replace: the set becomes exactly ids[0..n).  Otherwise ids are added to it (the table is rebuilt at twice
// the size when it would get more than half full).
int igs_idset_load(igs_engine* e, igs_idset* s, const uint64_t* ids, uint64_t n, bool replace) {
    cudaStream_t st = e->slot[0].stream;
    IGS_CUDA(e, cudaDeviceSynchronize());
    if (replace) igs_idset_free(s);
    if (n == 0) return IGS_OK;
    uint64_t* d_ids = nullptr;
    unsigned long long* d_added = nullptr;
    IGS_CUDA(e, cudaMalloc(&d_ids, n * 8 + 8));
    d_added = reinterpret_cast<unsigned long long*>(d_ids + n);
    int rc = IGS_OK;
    do {
        if (cudaMemcpyAsync(d_ids, ids, n * 8, cudaMemcpyHostToDevice, st) != cudaSuccess ||
            cudaMemsetAsync(d_added, 0, 8, st) != cudaSuccess) { rc = IGS_ERR_CUDA; break; }
        const uint64_t need = table_cap_for(s->count + n);       // upper bound: every id is new
        if (need > s->cap) {
            uint64_t* fresh = nullptr;
            if (cudaMalloc(&fresh, need * 8) != cudaSuccess) { rc = IGS_ERR_NOMEM; break; }
            cudaMemsetAsync(fresh, 0, need * 8, st);
            if (s->slots) {        // re-insert the old table's slots (empty ones are id 0 and are skipped)
                unsigned long long* d_tmp = nullptr;
                if (cudaMalloc(&d_tmp, 8) != cudaSuccess) { cudaFree(fresh); rc = IGS_ERR_NOMEM; break; }
                cudaMemsetAsync(d_tmp, 0, 8, st);
                k_idset_insert<<<(unsigned)((s->cap + kThreads - 1) / kThreads), kThreads, 0, st>>>(
                    reinterpret_cast<unsigned long long*>(fresh), need - 1, s->slots, s->cap, d_tmp);
                cudaStreamSynchronize(st);
                cudaFree(d_tmp);
                cudaFree(s->slots);
            }
            s->slots = fresh;
            s->cap = need;
        }
        k_idset_insert<<<(unsigned)((n + kThreads - 1) / kThreads), kThreads, 0, st>>>(
            reinterpret_cast<unsigned long long*>(s->slots), s->cap - 1, d_ids, n, d_added);
        unsigned long long added = 0;
        if (cudaGetLastError() != cudaSuccess ||
            cudaMemcpyAsync(&added, d_added, 8, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
            cudaStreamSynchronize(st) != cudaSuccess) { rc = IGS_ERR_CUDA; break; }
        s->count += added;
    } while (false);
    cudaFree(d_ids);
    if (rc == IGS_ERR_CUDA) igs_set_error(e, "blacklist upload failed: %s", cudaGetErrorString(cudaGetLastError()));
    if (rc == IGS_ERR_NOMEM) igs_set_error(e, "out of device memory for a blacklist table");
    return rc;
}

static igs_idset_view view_of(const igs_idset& s) {
    igs_idset_view v;
    v.slots = s.slots;
    v.mask = s.cap ? s.cap - 1 : 0;
    return v;
}

void igs_blacklist_views(const igs_engine* e, igs_idset_view out[3]) {
    for (int k = 0; k < 3; ++k) out[k] = view_of(e->bl[k]);
}

int igs_launch_blacklist_check(igs_engine* e, const igs_tx_ids* d_ids, uint64_t n, uint8_t* d_out,
                               cudaStream_t stream) {
    if (n == 0) return IGS_OK;
    k_blacklist_check<<<(unsigned)((n + kThreads - 1) / kThreads), kThreads, 0, stream>>>(
        view_of(e->bl[IGS_BL_DEVICE]), view_of(e->bl[IGS_BL_FINGERPRINT]), view_of(e->bl[IGS_BL_IP]), d_ids, n, d_out);
    IGS_LAUNCHED(e);
    return IGS_OK;
}

extern "C" uint64_t igs_hash_id(const char* s, size_t len) {
    if (s == nullptr || len == 0) return 0;
    uint64_t h = 0xcbf29ce484222325ull;                       // FNV-1a 64
    for (size_t i = 0; i < len; ++i) { h ^= (uint8_t)s[i]; h *= 0x100000001b3ull; }
    h = igs_idset_mix(h);
    return h ? h : 1;
}
