/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's pointnet2_batch algorithms.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library; it is the CHECKER, never the product path (the product is csrc/*.cu behind
 * include/apb200.h and fails loudly when the CUDA library is missing).
 *
 * Parity status: PINNED.  The reference ships no golden vectors for this path, so
 * this restatement is pinned against the reference's own CUDA kernels compiled unmodified from
 * /root/reference into oracle/_ref/libpointnet2_ref.so (oracle/Makefile) and executed on the B200
 * box: tests/test_parity_gpu.py::test_oracle_vs_compiled_reference_* compare them bit for bit,
 * and tests/golden/*.npz holds vectors generated from that run (tests/golden/make_golden.py).
 *
 * Every function cites the reference file:line it follows (paths relative to
 * /root/reference/openpoints/cpp/pointnet2_batch/src/).  Floating-point order: nvcc (-O2, default
 * --fmad=true) contracts  a*a + b*b + c*c  of the reference kernels into
 *      t = b*b (rounded) ; t = fma(a,a,t) ; t = fma(c,c,t)
 * (verified in the sm_100a SASS of the compiled reference: FMUL on the y term, then FFMA x, FFMA z;
 * same for three_interpolate: FMUL w1*p1, FFMA w0*p0, FFMA w2*p2).  Built with -ffp-contract=off
 * so that only the explicit fmaf() calls below fuse.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline float sqdist_ref(float ax, float ay, float az, float bx, float by, float bz) {
    /* sampling_gpu.cu:140, ball_query_gpu.cu:40, interpolate_gpu.cu:42 (+ SASS contraction order) */
    float dx = ax - bx, dy = ay - by, dz = az - bz;
    float t = dy * dy;
    t = fmaf(dx, dx, t);
    t = fmaf(dz, dz, t);
    return t;
}

/* cuda_utils.h:10-14 : largest power of two <= work_size, capped at 1024, computed through
 * double log() exactly like the reference host code (so rounding quirks are shared). */
int orc_opt_n_threads(int work_size) {
    const int pow_2 = (int)(log((double)work_size) / log(2.0));
    int v = 1 << pow_2;
    if (v > 1024) v = 1024;
    if (v < 1) v = 1;
    return v;
}

/* CUDA fminf/min(float,float) == FMNMX: returns the non-NaN operand if exactly one is NaN. */
static inline float fmin_cuda(float a, float b) { return fminf(a, b); }

/* sampling_gpu.cu:100-216 (kernel) + 218-260 (launcher).  Literal emulation: `bs` threads, thread
 * tid scans k = tid, tid+bs, ... with strict '>' from (best=-1, besti=0), then the shared-memory
 * tree with offsets bs/2 ... 1 in which the lower slot wins ties (sampling_gpu.cu:93-98).
 * temp: (B,N) scratch, caller pre-fills 1e10 (subsample.py:94); updated in place like the kernel. */
void orc_fps(int b, int n, int m, const float *xyz, float *temp, int32_t *idxs) {
    if (m <= 0) return;
    const int bs = orc_opt_n_threads(n);
    float *dists = (float *)malloc(sizeof(float) * bs);
    int *dists_i = (int *)malloc(sizeof(int) * bs);
    for (int bi = 0; bi < b; ++bi) {
        const float *ds = xyz + (size_t)bi * n * 3;
        float *tp = temp + (size_t)bi * n;
        int32_t *out = idxs + (size_t)bi * m;
        int old = 0;
        out[0] = 0;
        for (int j = 1; j < m; ++j) {
            const float x1 = ds[old * 3 + 0], y1 = ds[old * 3 + 1], z1 = ds[old * 3 + 2];
            for (int tid = 0; tid < bs; ++tid) {
                int besti = 0;
                float best = -1.f;
                for (int k = tid; k < n; k += bs) {
                    float d = sqdist_ref(ds[k * 3 + 0], ds[k * 3 + 1], ds[k * 3 + 2], x1, y1, z1);
                    float d2 = fmin_cuda(d, tp[k]);
                    tp[k] = d2;
                    besti = d2 > best ? k : besti;
                    best = d2 > best ? d2 : best;
                }
                dists[tid] = best;
                dists_i[tid] = besti;
            }
            for (int off = bs / 2; off >= 1; off >>= 1) {
                for (int tid = 0; tid < off; ++tid) {
                    const float v1 = dists[tid], v2 = dists[tid + off];
                    const int i1 = dists_i[tid], i2 = dists_i[tid + off];
                    dists[tid] = fmaxf(v1, v2);
                    dists_i[tid] = v2 > v1 ? i2 : i1;
                }
            }
            old = dists_i[0];
            out[j] = old;
        }
    }
    free(dists);
    free(dists_i);
}

/* ball_query_gpu.cu:15-51.  idx (B,M,K) is pre-zeroed by the caller (group.py:194); a query with
 * no hit keeps its zeros. */
void orc_ball_query(int b, int n, int m, float radius, int nsample, const float *new_xyz,
                    const float *xyz, int32_t *idx) {
    const float radius2 = radius * radius;
    for (int bi = 0; bi < b; ++bi)
        for (int j = 0; j < m; ++j) {
            const float *q = new_xyz + ((size_t)bi * m + j) * 3;
            const float *p = xyz + (size_t)bi * n * 3;
            int32_t *o = idx + ((size_t)bi * m + j) * nsample;
            int cnt = 0;
            for (int k = 0; k < n; ++k) {
                float d2 = sqdist_ref(q[0], q[1], q[2], p[k * 3 + 0], p[k * 3 + 1], p[k * 3 + 2]);
                if (d2 < radius2) {
                    if (cnt == 0)
                        for (int l = 0; l < nsample; ++l) o[l] = k;
                    o[cnt] = k;
                    ++cnt;
                    if (cnt >= nsample) break;
                }
            }
        }
}

/* group_points_gpu.cu:53-72 */
void orc_group(int b, int c, int n, int m, int k, const float *points, const int32_t *idx, float *out) {
    for (int bi = 0; bi < b; ++bi)
        for (int ci = 0; ci < c; ++ci) {
            const float *row = points + ((size_t)bi * c + ci) * n;
            float *o = out + ((size_t)bi * c + ci) * m * k;
            const int32_t *ix = idx + (size_t)bi * m * k;
            for (size_t p = 0; p < (size_t)m * k; ++p) o[p] = row[ix[p]];
        }
}

/* group_points_gpu.cu:14-31.  grad_points pre-zeroed by the caller (group.py:111).  The reference
 * sums with fp32 atomics in an unspecified order; this restatement sums in ascending position
 * order with a double accumulator per target so it is the better-conditioned comparand. */
void orc_group_grad(int b, int c, int n, int m, int k, const float *grad_out, const int32_t *idx,
                    float *grad_points) {
    double *acc = (double *)malloc(sizeof(double) * n);
    for (int bi = 0; bi < b; ++bi)
        for (int ci = 0; ci < c; ++ci) {
            float *row = grad_points + ((size_t)bi * c + ci) * n;
            const float *g = grad_out + ((size_t)bi * c + ci) * m * k;
            const int32_t *ix = idx + (size_t)bi * m * k;
            for (int i = 0; i < n; ++i) acc[i] = row[i];
            for (size_t p = 0; p < (size_t)m * k; ++p) acc[ix[p]] += g[p];
            for (int i = 0; i < n; ++i) row[i] = (float)acc[i];
        }
    free(acc);
}

/* sampling_gpu.cu:15-31 / 53-70 : the K=1 case of grouping */
void orc_gather(int b, int c, int n, int m, const float *points, const int32_t *idx, float *out) {
    orc_group(b, c, n, m, 1, points, idx, out);
}
void orc_gather_grad(int b, int c, int n, int m, const float *grad_out, const int32_t *idx, float *grad_points) {
    orc_group_grad(b, c, n, m, 1, grad_out, idx, grad_points);
}

/* interpolate_gpu.cu:16-59.  The running bests are double initialised to 1e40 and compared with
 * the float distance (strict '<'); the stores convert back to float (1e40 -> +inf). */
void orc_three_nn(int b, int n, int m, const float *unknown, const float *known, float *dist2, int32_t *idx) {
    for (int bi = 0; bi < b; ++bi)
        for (int i = 0; i < n; ++i) {
            const float *u = unknown + ((size_t)bi * n + i) * 3;
            const float *kn = known + (size_t)bi * m * 3;
            double best1 = 1e40, best2 = 1e40, best3 = 1e40;
            int besti1 = 0, besti2 = 0, besti3 = 0;
            for (int k = 0; k < m; ++k) {
                float d = sqdist_ref(u[0], u[1], u[2], kn[k * 3 + 0], kn[k * 3 + 1], kn[k * 3 + 2]);
                if (d < best1) {
                    best3 = best2; besti3 = besti2;
                    best2 = best1; besti2 = besti1;
                    best1 = d; besti1 = k;
                } else if (d < best2) {
                    best3 = best2; besti3 = besti2;
                    best2 = d; besti2 = k;
                } else if (d < best3) {
                    best3 = d; besti3 = k;
                }
            }
            float *od = dist2 + ((size_t)bi * n + i) * 3;
            int32_t *oi = idx + ((size_t)bi * n + i) * 3;
            od[0] = (float)best1; od[1] = (float)best2; od[2] = (float)best3;
            oi[0] = besti1; oi[1] = besti2; oi[2] = besti3;
        }
}

/* interpolate_gpu.cu:84-104 (+ SASS order: w1*p1, then fma w0*p0, then fma w2*p2) */
void orc_three_interp(int b, int c, int m, int n, const float *points, const int32_t *idx,
                      const float *weight, float *out) {
    for (int bi = 0; bi < b; ++bi)
        for (int ci = 0; ci < c; ++ci) {
            const float *row = points + ((size_t)bi * c + ci) * m;
            float *o = out + ((size_t)bi * c + ci) * n;
            for (int i = 0; i < n; ++i) {
                const float *w = weight + ((size_t)bi * n + i) * 3;
                const int32_t *ix = idx + ((size_t)bi * n + i) * 3;
                float t = w[1] * row[ix[1]];
                t = fmaf(w[0], row[ix[0]], t);
                t = fmaf(w[2], row[ix[2]], t);
                o[i] = t;
            }
        }
}

/* interpolate_gpu.cu:127-149 ; grad_points (B,C,M) pre-zeroed (upsampling.py:82).  Double
 * accumulation in ascending (i, j) order, see orc_group_grad. */
void orc_three_interp_grad(int b, int c, int n, int m, const float *grad_out, const int32_t *idx,
                           const float *weight, float *grad_points) {
    double *acc = (double *)malloc(sizeof(double) * m);
    for (int bi = 0; bi < b; ++bi)
        for (int ci = 0; ci < c; ++ci) {
            float *row = grad_points + ((size_t)bi * c + ci) * m;
            const float *g = grad_out + ((size_t)bi * c + ci) * n;
            for (int i = 0; i < m; ++i) acc[i] = row[i];
            for (int i = 0; i < n; ++i) {
                const float *w = weight + ((size_t)bi * n + i) * 3;
                const int32_t *ix = idx + ((size_t)bi * n + i) * 3;
                for (int j = 0; j < 3; ++j) acc[ix[j]] += (double)(g[i] * w[j]);
            }
            for (int i = 0; i < m; ++i) row[i] = (float)acc[i];
        }
    free(acc);
}

/* kNN: group.py:12-28 (torch.cdist + topk(largest=False)).  The arithmetic lives in PyTorch
 * (third-party, reference pins torch==1.12.1+cu113, README.md:26); topk tie order is unspecified
 * there, so this oracle DEFINES the order as (squared distance, then index) ascending, using the
 * same fp32 distance expression as the other ops.  Parity for kNN vs torch is "unpinned":
 * tests check set equality away from ties.  dist2 (B,M,K) optional. */
void orc_knn(int b, int n, int m, int k, const float *new_xyz, const float *xyz, int32_t *idx, float *dist2) {
    float *bd = (float *)malloc(sizeof(float) * k);
    int *bi_ = (int *)malloc(sizeof(int) * k);
    for (int bi = 0; bi < b; ++bi)
        for (int j = 0; j < m; ++j) {
            const float *q = new_xyz + ((size_t)bi * m + j) * 3;
            const float *p = xyz + (size_t)bi * n * 3;
            int cnt = 0;
            for (int t = 0; t < n; ++t) {
                float d = sqdist_ref(q[0], q[1], q[2], p[t * 3 + 0], p[t * 3 + 1], p[t * 3 + 2]);
                if (cnt == k && !(d < bd[k - 1])) continue;
                int pos = cnt < k ? cnt : k - 1;
                while (pos > 0 && d < bd[pos - 1]) { bd[pos] = bd[pos - 1]; bi_[pos] = bi_[pos - 1]; --pos; }
                bd[pos] = d; bi_[pos] = t;
                if (cnt < k) ++cnt;
            }
            for (int t = 0; t < k; ++t) {
                idx[((size_t)bi * m + j) * k + t] = t < cnt ? bi_[t] : 0;
                if (dist2) dist2[((size_t)bi * m + j) * k + t] = t < cnt ? bd[t] : INFINITY;
            }
        }
    free(bd);
    free(bi_);
}
