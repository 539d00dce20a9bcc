// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// extern "C" shim over the UNMODIFIED reference CUDA launchers of
// openpoints/cpp/pointnet2_batch (sampling_gpu.cu, ball_query_gpu.cu,
// group_points_gpu.cu, interpolate_gpu.cu).  The reference sources are compiled
// where they lie under /root/reference by oracle/Makefile; nothing is copied.
// The launchers are plain C++ functions over raw device pointers
// (prototypes: src/sampling_gpu.h:14-30, ball_query_gpu.h:13-14,
// group_points_gpu.h:14-21, interpolate_gpu.h:14-33), they launch on the legacy
// default stream, so callers must synchronise around them.
#include <cuda_runtime.h>

void gather_points_kernel_launcher_fast(int b, int c, int n, int npoints,
    const float *points, const int *idx, float *out);
void gather_points_grad_kernel_launcher_fast(int b, int c, int n, int npoints,
    const float *grad_out, const int *idx, float *grad_points);
void furthest_point_sampling_kernel_launcher(int b, int n, int m,
    const float *dataset, float *temp, int *idxs);
void ball_query_kernel_launcher_fast(int b, int n, int m, float radius, int nsample,
    const float *new_xyz, const float *xyz, int *idx);
void group_points_kernel_launcher_fast(int b, int c, int n, int npoints, int nsample,
    const float *points, const int *idx, float *out);
void group_points_grad_kernel_launcher_fast(int b, int c, int n, int npoints, int nsample,
    const float *grad_out, const int *idx, float *grad_points);
void three_nn_kernel_launcher_fast(int b, int n, int m, const float *unknown,
    const float *known, float *dist2, int *idx);
void three_interpolate_kernel_launcher_fast(int b, int c, int m, int n,
    const float *points, const int *idx, const float *weight, float *out);
void three_interpolate_grad_kernel_launcher_fast(int b, int c, int n, int m,
    const float *grad_out, const int *idx, const float *weight, float *grad_points);

extern "C" {
void ref_gather(int b, int c, int n, int m, const float *p, const int *i, float *o) {
    gather_points_kernel_launcher_fast(b, c, n, m, p, i, o); }
void ref_gather_grad(int b, int c, int n, int m, const float *g, const int *i, float *gp) {
    gather_points_grad_kernel_launcher_fast(b, c, n, m, g, i, gp); }
void ref_fps(int b, int n, int m, const float *xyz, float *temp, int *idx) {
    furthest_point_sampling_kernel_launcher(b, n, m, xyz, temp, idx); }
void ref_ball_query(int b, int n, int m, float r, int k, const float *new_xyz, const float *xyz, int *idx) {
    ball_query_kernel_launcher_fast(b, n, m, r, k, new_xyz, xyz, idx); }
void ref_group(int b, int c, int n, int m, int k, const float *p, const int *i, float *o) {
    group_points_kernel_launcher_fast(b, c, n, m, k, p, i, o); }
void ref_group_grad(int b, int c, int n, int m, int k, const float *g, const int *i, float *gp) {
    group_points_grad_kernel_launcher_fast(b, c, n, m, k, g, i, gp); }
void ref_three_nn(int b, int n, int m, const float *u, const float *kn, float *d2, int *idx) {
    three_nn_kernel_launcher_fast(b, n, m, u, kn, d2, idx); }
void ref_three_interp(int b, int c, int m, int n, const float *p, const int *i, const float *w, float *o) {
    three_interpolate_kernel_launcher_fast(b, c, m, n, p, i, w, o); }
void ref_three_interp_grad(int b, int c, int n, int m, const float *g, const int *i, const float *w, float *gp) {
    three_interpolate_grad_kernel_launcher_fast(b, c, n, m, g, i, w, gp); }
int ref_sync() { return (int)cudaDeviceSynchronize(); }
}
