// k_ltv.cu — placeholder until the tcgen05 MLP lands.
#include "igs_internal.cuh"

int igs_mlp_upload(igs_engine* e, const igs_mlp_desc*) {
    igs_set_error(e, "LTV MLP: not built into this library yet");
    return IGS_ERR_STATE;
}
void igs_mlp_free(igs_engine*) {}
int igs_launch_ltv(igs_engine* e, const uint16_t*, const igs_ltv_row*, uint64_t, igs_ltv_result*, float*,
                   cudaStream_t) {
    igs_set_error(e, "LTV MLP: not built into this library yet");
    return IGS_ERR_STATE;
}
