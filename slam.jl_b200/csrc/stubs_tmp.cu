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
int slam_shard_unique_id(char *) NI
int slam_shard_create(int, int, int, const char *, int, slam_shard_t **) NI
int slam_shard_destroy(slam_shard_t *) NI
int slam_shard_plan(int, int, int, int *, int *) NI
int slam_shard_synth_state(slam_shard_t *, int, double, uint64_t) NI
int slam_shard_n(slam_shard_t *) NI
int slam_shard_get_pose(slam_shard_t *, double *) NI
int slam_shard_get_x(slam_shard_t *, double *) NI
int slam_shard_get_cov_block(slam_shard_t *, int, int, int, int, double *, int) NI
int slam_shard_predict(slam_shard_t *, double, double, double, const double *, double) NI
int slam_shard_associate(slam_shard_t *, const double *, int, const double *, double, double, double *, int64_t *, int *, double *, int *) NI
int slam_shard_update(slam_shard_t *, const double *, int, const double *, const int64_t *) NI
int slam_shard_update_resident(slam_shard_t *, const double *, int, const double *, const int64_t *) NI
int slam_shard_sync(slam_shard_t *) NI
void *slam_shard_stream(slam_shard_t *) { return nullptr; }
long long slam_shard_launch_count(slam_shard_t *) { return 0; }
}
