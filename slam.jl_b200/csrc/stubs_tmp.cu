// temporary: entry points not implemented yet
#include "common.cuh"
#define NI { slam_set_error("%s: not implemented yet", __func__); return SLAM_ERR_ARG; }
extern "C" {
int slam_ens_create(int, const double *, int, const double *, int, int, uint64_t, uint32_t, int, slam_ens_t **) NI
int slam_ens_destroy(slam_ens_t *) NI
int slam_ens_steps(slam_ens_t *, int *, int *) NI
int slam_ens_run(slam_ens_t *) NI
int slam_ens_sync(slam_ens_t *) NI
int slam_ens_get_results(slam_ens_t *, double *, int *, int *) NI
int slam_ens_keep_tracks(slam_ens_t *, int) NI
int slam_ens_get_track(slam_ens_t *, int, double *) NI
int slam_ens_get_state(slam_ens_t *, int, double *, double *, int) NI
int slam_ens_set_noise(slam_ens_t *, const double *, const double *, int) NI
void *slam_ens_stream(slam_ens_t *) { return nullptr; }
long long slam_ens_launch_count(slam_ens_t *) { return 0; }
}
