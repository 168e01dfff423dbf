// k_ltv_heur.cu — the reference's model-free LTV arithmetic on the device (sm_100a), one player per thread.
//   k_ltv_predictor : LTVPredictor.Predict (prediction/ltv.go:113-151) with calculateLTV (:155-178) in the model
//                     slot — what LTVPredictor.BatchPredict (:385-398) computes today, sequentially.
//   k_ltv_mock      : ml.LTVModel.Predict (ml/onnx_model.go:405-432) — "session: nil // Use mock for now".
// Pure fp64 streaming work: 80 B in + 32 (+8) B out per player, resp. 68 B in + 8 B out.  HBM-bound.
#include "igs_internal.cuh"
#include "ltv_math.cuh"

namespace {

constexpr int kHeurThreads = 256;

__global__ void __launch_bounds__(kHeurThreads)
k_ltv_predictor(const igs_ltv_row* __restrict__ rows, unsigned long long n, igs_ltv_result* __restrict__ out,
                double* __restrict__ raw) {
    const unsigned long long i = (unsigned long long)blockIdx.x * kHeurThreads + threadIdx.x;
    if (i >= n) return;
    union __align__(16) { igs_ltv_row r; uint4 q[sizeof(igs_ltv_row) / 16]; } in;
    const uint4* src = reinterpret_cast<const uint4*>(rows + i);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(igs_ltv_row) / 16); ++k) in.q[k] = __ldg(src + k);
    const double ltv = ltv_calculate(in.r);
    union __align__(16) { igs_ltv_result r; uint4 q[2]; } res;
    ltv_epilogue(in.r, ltv, &res.r);
    uint4* dst = reinterpret_cast<uint4*>(out + i);
    dst[0] = res.q[0];
    dst[1] = res.q[1];
    if (raw) raw[i] = ltv;
}

__global__ void __launch_bounds__(kHeurThreads)
k_ltv_mock(const igs_ltv_features* __restrict__ f, unsigned long long n, double* __restrict__ out) {
    const unsigned long long i = (unsigned long long)blockIdx.x * kHeurThreads + threadIdx.x;
    if (i >= n) return;
    igs_ltv_features r;
    const float* src = reinterpret_cast<const float*>(f + i);          // 68-byte rows: 4-byte aligned only
    float* dst = reinterpret_cast<float*>(&r);
#pragma unroll
    for (int k = 0; k < (int)(sizeof(igs_ltv_features) / 4); ++k) dst[k] = __ldg(src + k);
    out[i] = ltv_mock_predict(r);
}

int launch_predictor(igs_engine* e, const igs_ltv_row* d_rows, uint64_t n, igs_ltv_result* d_out, double* d_raw,
                     cudaStream_t st) {
    if (n == 0) return IGS_OK;
    const uint64_t nb = (n + kHeurThreads - 1) / kHeurThreads;
    IGS_REQUIRE(e, nb <= 0x7FFFFFFFull, "too many rows for one launch");
    k_ltv_predictor<<<(unsigned)nb, kHeurThreads, 0, st>>>(d_rows, n, d_out, d_raw);
    IGS_LAUNCHED(e);
    return IGS_OK;
}

int launch_mock(igs_engine* e, const igs_ltv_features* d_f, uint64_t n, double* d_out, cudaStream_t st) {
    if (n == 0) return IGS_OK;
    const uint64_t nb = (n + kHeurThreads - 1) / kHeurThreads;
    IGS_REQUIRE(e, nb <= 0x7FFFFFFFull, "too many rows for one launch");
    k_ltv_mock<<<(unsigned)nb, kHeurThreads, 0, st>>>(d_f, n, d_out);
    IGS_LAUNCHED(e);
    return IGS_OK;
}

int set_device(igs_engine* e) {
    if (!e) {
        igs_set_error(nullptr, "engine handle is null");
        return IGS_ERR_INVALID;
    }
    IGS_CUDA(e, cudaSetDevice(e->device));
    return IGS_OK;
}

}  // namespace

extern "C" {

int igs_ltv_predict_batch_device(igs_engine* e, const igs_ltv_row* d_rows, uint64_t n, igs_ltv_result* d_out,
                                 double* d_ltv_raw, void* stream) {
    int rc = set_device(e);
    if (rc != IGS_OK) return rc;
    IGS_REQUIRE(e, n == 0 || (d_rows && d_out), "rows / out is null");
    return launch_predictor(e, d_rows, n, d_out, d_ltv_raw, (cudaStream_t)stream);
}

int igs_ltv_predict_batch(igs_engine* e, const igs_ltv_row* rows, uint64_t n, igs_ltv_result* out, double* ltv_raw) {
    int rc = set_device(e);
    if (rc != IGS_OK) return rc;
    IGS_REQUIRE(e, n == 0 || (rows && out), "rows / out is null");
    if (n == 0) return IGS_OK;
    igs_slot& sl = e->slot[0];
    if ((rc = igs_reserve(e, &sl.in1, n * sizeof(igs_ltv_row))) != IGS_OK) return rc;
    if ((rc = igs_reserve(e, &sl.out0, n * sizeof(igs_ltv_result))) != IGS_OK) return rc;
    if (ltv_raw && (rc = igs_reserve(e, &sl.aux0, n * sizeof(double))) != IGS_OK) return rc;
    IGS_CUDA(e, cudaMemcpyAsync(sl.in1.p, rows, n * sizeof(igs_ltv_row), cudaMemcpyHostToDevice, sl.stream));
    rc = launch_predictor(e, (const igs_ltv_row*)sl.in1.p, n, (igs_ltv_result*)sl.out0.p,
                          ltv_raw ? (double*)sl.aux0.p : nullptr, sl.stream);
    if (rc == IGS_OK) {
        cudaMemcpyAsync(out, sl.out0.p, n * sizeof(igs_ltv_result), cudaMemcpyDeviceToHost, sl.stream);
        if (ltv_raw) cudaMemcpyAsync(ltv_raw, sl.aux0.p, n * sizeof(double), cudaMemcpyDeviceToHost, sl.stream);
    }
    IGS_CUDA(e, cudaStreamSynchronize(sl.stream));      // also on failure: nothing may still write into `out`
    return rc;
}

int igs_ltv_mock_predict_device(igs_engine* e, const igs_ltv_features* d_f, uint64_t n, double* d_ltv_out,
                                void* stream) {
    int rc = set_device(e);
    if (rc != IGS_OK) return rc;
    IGS_REQUIRE(e, n == 0 || (d_f && d_ltv_out), "features / out is null");
    return launch_mock(e, d_f, n, d_ltv_out, (cudaStream_t)stream);
}

int igs_ltv_mock_predict(igs_engine* e, const igs_ltv_features* f, uint64_t n, double* ltv_out) {
    int rc = set_device(e);
    if (rc != IGS_OK) return rc;
    IGS_REQUIRE(e, n == 0 || (f && ltv_out), "features / out is null");
    if (n == 0) return IGS_OK;
    igs_slot& sl = e->slot[0];
    if ((rc = igs_reserve(e, &sl.in1, n * sizeof(igs_ltv_features))) != IGS_OK) return rc;
    if ((rc = igs_reserve(e, &sl.aux0, n * sizeof(double))) != IGS_OK) return rc;
    IGS_CUDA(e, cudaMemcpyAsync(sl.in1.p, f, n * sizeof(igs_ltv_features), cudaMemcpyHostToDevice, sl.stream));
    rc = launch_mock(e, (const igs_ltv_features*)sl.in1.p, n, (double*)sl.aux0.p, sl.stream);
    if (rc == IGS_OK) cudaMemcpyAsync(ltv_out, sl.aux0.p, n * sizeof(double), cudaMemcpyDeviceToHost, sl.stream);
    IGS_CUDA(e, cudaStreamSynchronize(sl.stream));
    return rc;
}

}  // extern "C"
