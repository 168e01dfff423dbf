// k_velocity.cu — placeholder until the event index lands (next commit).
#include "igs_internal.cuh"

int igs_build_event_index(igs_engine* e, const igs_event_cols*, uint64_t, uint32_t) {
    igs_set_error(e, "event index: not built into this library yet");
    return IGS_ERR_STATE;
}
void igs_free_event_index(igs_engine*) {}
int igs_launch_velocity(igs_engine* e, const igs_query*, uint64_t, igs_velocity*, cudaStream_t) {
    igs_set_error(e, "velocity: not built into this library yet");
    return IGS_ERR_STATE;
}
int igs_launch_score_batch(igs_engine* e, const igs_config*, const igs_tx*, uint64_t, igs_result*, float*,
                           igs_velocity*, cudaStream_t) {
    igs_set_error(e, "fused path: not built into this library yet");
    return IGS_ERR_STATE;
}
