// BatchNorm (training mode, batch statistics) (+ReLU) and BatchNorm + max-pool over the K neighbours, for the
// channel-major (B, C, P) activations of the shared MLP.  HBM-bound streaming kernels.
//
// They replace, behind the same nn.BatchNorm{1,2}d parameters (openpoints/models/layers/norm.py:57-71 'bn1d'/'bn2d',
// conv.py:24-103 conv-norm-act blocks) and the max reduction of SetAbstraction / LocalAggregation
// (pointnext.py:134, :74-77), what torch runs as cudnn/native batch_norm (+threshold) (+max over the last dim):
//   forward  BN(+ReLU):  one read pass for the statistics, one read+write pass to apply      (torch: same + ReLU pass)
//   backward BN(+ReLU):  one read pass (sum dy, sum dy*xhat), one read+write pass            (torch native: ~5x slower)
//   forward  BN -> max_K: ONE read pass: per-channel sum / sum of squares AND per-(b,c,m) max, min, arg over K;
//                         BN is monotone per channel, so max_k BN(y) = BN(max_k y) if gamma >= 0 else BN(min_k y):
//                         the normalised (B,C,M,K) tensor is never written;
//   backward BN -> max_K: dy = gsel * [k == arg] + a[c] + b[c] * y   (the batch-statistics terms of the BatchNorm
//                         backward are affine in y), one read + one write pass.
// Statistics are accumulated per CTA in fp32 over <= 4096 elements and across CTAs in fp64 (atomicAdd double).
#include "apb_common.cuh"

namespace {

constexpr int NT = 256;
constexpr int CHUNK = 4096;  // elements of one (b, c) row handled by a CTA

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// block-wide sum of two values; result valid in thread 0
__device__ __forceinline__ void block_sum2(float &a, float &b) {
    __shared__ float red[2][NT / 32];
    a = warp_sum(a);
    b = warp_sum(b);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { red[0][warp] = a; red[1][warp] = b; }
    __syncthreads();
    if (warp == 0) {
        a = lane < NT / 32 ? red[0][lane] : 0.f;
        b = lane < NT / 32 ? red[1][lane] : 0.f;
        a = warp_sum(a);
        b = warp_sum(b);
    }
}

// stats[c] += (sum x, sum x^2) over this CTA's chunk of row (b, c)
__global__ void __launch_bounds__(NT) bn_stats_kernel(const float *__restrict__ x, double *__restrict__ stats, int C,
                                                      long long P) {
    const int c = blockIdx.y, b = blockIdx.z;
    const float *row = x + ((size_t)b * C + c) * P;
    const long long p0 = (long long)blockIdx.x * CHUNK;
    const long long p1 = min(P, p0 + CHUNK);
    float s = 0.f, q = 0.f;
    if ((P & 3) == 0 && (reinterpret_cast<uintptr_t>(row) & 15) == 0) {
        for (long long p = p0 + threadIdx.x * 4; p < p1; p += NT * 4) {
            const float4 v = apb_ld_nc_f4(row + p);
            s += (v.x + v.y) + (v.z + v.w);
            q += (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w);
        }
    } else {
        for (long long p = p0 + threadIdx.x; p < p1; p += NT) {
            const float v = __ldg(row + p);
            s += v;
            q += v * v;
        }
    }
    block_sum2(s, q);
    if (threadIdx.x == 0) {
        atomicAdd(stats + 2 * c, (double)s);
        atomicAdd(stats + 2 * c + 1, (double)q);
    }
}

struct BnApplyParams {
    const float *x;
    float *y;
    const double *stats;
    const float *gamma, *beta;
    float eps;
    double inv_count;  // 1 / (elements per channel over all ranks)
    double unbias;     // n / (n - 1)
    int relu;
    float *save_mean, *save_rstd, *running_mean, *running_var;  // running_* may be null
    float momentum;
    int C;
    long long P;
};

__global__ void __launch_bounds__(NT) bn_apply_kernel(const BnApplyParams p) {
    const int c = blockIdx.y, b = blockIdx.z;
    const double mean_d = p.stats[2 * c] * p.inv_count;
    double var_d = p.stats[2 * c + 1] * p.inv_count - mean_d * mean_d;
    if (var_d < 0.0) var_d = 0.0;
    const float mean = (float)mean_d;
    const float rstd = (float)(1.0 / sqrt(var_d + (double)p.eps));
    const float scale = p.gamma[c] * rstd;
    const float shift = p.beta[c] - mean * scale;
    if (blockIdx.x == 0 && b == 0 && threadIdx.x == 0) {
        p.save_mean[c] = mean;
        p.save_rstd[c] = rstd;
        if (p.running_mean) {
            p.running_mean[c] = (1.f - p.momentum) * p.running_mean[c] + p.momentum * mean;
            p.running_var[c] = (1.f - p.momentum) * p.running_var[c] + p.momentum * (float)(var_d * p.unbias);
        }
    }
    const float *row = p.x + ((size_t)b * p.C + c) * p.P;
    float *out = p.y + ((size_t)b * p.C + c) * p.P;
    const long long p0 = (long long)blockIdx.x * CHUNK;
    const long long p1 = min(p.P, p0 + CHUNK);
    if ((p.P & 3) == 0 && ((reinterpret_cast<uintptr_t>(row) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
        for (long long i = p0 + threadIdx.x * 4; i < p1; i += NT * 4) {
            float4 v = apb_ld_nc_f4(row + i);
            v.x = fmaf(v.x, scale, shift); v.y = fmaf(v.y, scale, shift);
            v.z = fmaf(v.z, scale, shift); v.w = fmaf(v.w, scale, shift);
            if (p.relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
            *reinterpret_cast<float4 *>(out + i) = v;
        }
    } else {
        for (long long i = p0 + threadIdx.x; i < p1; i += NT) {
            float v = fmaf(__ldg(row + i), scale, shift);
            out[i] = p.relu ? fmaxf(v, 0.f) : v;
        }
    }
}

// stats2[c] += (sum dy', sum dy' * xhat), dy' = dy * [y > 0] when `y` (the ReLU output) is given
__global__ void __launch_bounds__(NT) bn_bwd_reduce_kernel(const float *__restrict__ x, const float *__restrict__ dy,
                                                           const float *__restrict__ y,
                                                           const float *__restrict__ save_mean,
                                                           const float *__restrict__ save_rstd,
                                                           double *__restrict__ stats2, int C, long long P) {
    const int c = blockIdx.y, b = blockIdx.z;
    const size_t off = ((size_t)b * C + c) * P;
    const float mean = save_mean[c], rstd = save_rstd[c];
    const long long p0 = (long long)blockIdx.x * CHUNK;
    const long long p1 = min(P, p0 + CHUNK);
    float s = 0.f, q = 0.f;
    const bool vec = (P & 3) == 0 && (((reinterpret_cast<uintptr_t>(x + off) | reinterpret_cast<uintptr_t>(dy + off) |
                                        (y ? reinterpret_cast<uintptr_t>(y + off) : 0)) & 15) == 0);
    if (vec) {
        for (long long i = p0 + threadIdx.x * 4; i < p1; i += NT * 4) {
            const float4 xv = apb_ld_nc_f4(x + off + i);
            float4 g = apb_ld_nc_f4(dy + off + i);
            if (y) {
                const float4 yv = apb_ld_nc_f4(y + off + i);
                g.x = yv.x > 0.f ? g.x : 0.f; g.y = yv.y > 0.f ? g.y : 0.f;
                g.z = yv.z > 0.f ? g.z : 0.f; g.w = yv.w > 0.f ? g.w : 0.f;
            }
            s += (g.x + g.y) + (g.z + g.w);
            q += (g.x * (xv.x - mean) + g.y * (xv.y - mean)) + (g.z * (xv.z - mean) + g.w * (xv.w - mean));
        }
    } else {
        for (long long i = p0 + threadIdx.x; i < p1; i += NT) {
            float g = __ldg(dy + off + i);
            if (y && !(__ldg(y + off + i) > 0.f)) g = 0.f;
            s += g;
            q += g * (__ldg(x + off + i) - mean);
        }
    }
    q *= rstd;
    block_sum2(s, q);
    if (threadIdx.x == 0) {
        atomicAdd(stats2 + 2 * c, (double)s);
        atomicAdd(stats2 + 2 * c + 1, (double)q);
    }
}

// dx = gamma * rstd * (dy' - sum(dy')/n - xhat * sum(dy' xhat)/n)
__global__ void __launch_bounds__(NT) bn_bwd_apply_kernel(const float *__restrict__ x, const float *__restrict__ dy,
                                                          const float *__restrict__ y,
                                                          const float *__restrict__ save_mean,
                                                          const float *__restrict__ save_rstd,
                                                          const float *__restrict__ gamma,
                                                          const double *__restrict__ stats2, double inv_count,
                                                          float *__restrict__ dx, int C, long long P) {
    const int c = blockIdx.y, b = blockIdx.z;
    const size_t off = ((size_t)b * C + c) * P;
    const float mean = save_mean[c], rstd = save_rstd[c];
    const float k0 = gamma[c] * rstd;
    const float m1 = (float)(stats2[2 * c] * inv_count);            // mean of dy'
    const float m2 = (float)(stats2[2 * c + 1] * inv_count) * rstd; // mean of dy' xhat, times rstd
    // dx = k0 * (g - m1 - (x - mean) * m2)
    const long long p0 = (long long)blockIdx.x * CHUNK;
    const long long p1 = min(P, p0 + CHUNK);
    const bool vec = (P & 3) == 0 && (((reinterpret_cast<uintptr_t>(x + off) | reinterpret_cast<uintptr_t>(dy + off) |
                                        reinterpret_cast<uintptr_t>(dx + off) |
                                        (y ? reinterpret_cast<uintptr_t>(y + off) : 0)) & 15) == 0);
    if (vec) {
        for (long long i = p0 + threadIdx.x * 4; i < p1; i += NT * 4) {
            const float4 xv = apb_ld_nc_f4(x + off + i);
            float4 g = apb_ld_nc_f4(dy + off + i);
            if (y) {
                const float4 yv = apb_ld_nc_f4(y + off + i);
                g.x = yv.x > 0.f ? g.x : 0.f; g.y = yv.y > 0.f ? g.y : 0.f;
                g.z = yv.z > 0.f ? g.z : 0.f; g.w = yv.w > 0.f ? g.w : 0.f;
            }
            float4 o;
            o.x = k0 * (g.x - m1 - (xv.x - mean) * m2); o.y = k0 * (g.y - m1 - (xv.y - mean) * m2);
            o.z = k0 * (g.z - m1 - (xv.z - mean) * m2); o.w = k0 * (g.w - m1 - (xv.w - mean) * m2);
            *reinterpret_cast<float4 *>(dx + off + i) = o;
        }
    } else {
        for (long long i = p0 + threadIdx.x; i < p1; i += NT) {
            float g = __ldg(dy + off + i);
            if (y && !(__ldg(y + off + i) > 0.f)) g = 0.f;
            dx[off + i] = k0 * (g - m1 - (__ldg(x + off + i) - mean) * m2);
        }
    }
}

// ---- short rows (P <= kSmallP, e.g. the imitator's anchor head: (32, 1024, 4)) -----------------------------------------
// The kernels above give every (b, c) row its own CTA; with a handful of elements per row that is tens of thousands
// of one-thread CTAs (35-50 us for 512 KB).  Here a THREAD owns a (b, c) row: thread <-> channel, blockIdx.y = b, so a
// warp reads 32*P contiguous floats.  Same math; per-channel sums still meet in the fp64 atomics.
constexpr int kSmallP = 64;

__global__ void __launch_bounds__(NT) bn_stats_small_kernel(const float *__restrict__ x, double *__restrict__ stats,
                                                            int C, int P) {
    const int c = blockIdx.x * NT + threadIdx.x, b = blockIdx.y;
    if (c >= C) return;
    const float *row = x + ((size_t)b * C + c) * P;
    float s = 0.f, q = 0.f;
    for (int i = 0; i < P; ++i) {
        const float v = __ldg(row + i);
        s += v;
        q += v * v;
    }
    atomicAdd(stats + 2 * c, (double)s);
    atomicAdd(stats + 2 * c + 1, (double)q);
}

__global__ void __launch_bounds__(NT) bn_apply_small_kernel(const BnApplyParams p) {
    const int c = blockIdx.x * NT + threadIdx.x, b = blockIdx.y;
    if (c >= p.C) return;
    const double mean_d = p.stats[2 * c] * p.inv_count;
    double var_d = p.stats[2 * c + 1] * p.inv_count - mean_d * mean_d;
    if (var_d < 0.0) var_d = 0.0;
    const float mean = (float)mean_d;
    const float rstd = (float)(1.0 / sqrt(var_d + (double)p.eps));
    const float scale = p.gamma[c] * rstd;
    const float shift = p.beta[c] - mean * scale;
    if (b == 0) {
        p.save_mean[c] = mean;
        p.save_rstd[c] = rstd;
        if (p.running_mean) {
            p.running_mean[c] = (1.f - p.momentum) * p.running_mean[c] + p.momentum * mean;
            p.running_var[c] = (1.f - p.momentum) * p.running_var[c] + p.momentum * (float)(var_d * p.unbias);
        }
    }
    const int P = (int)p.P;
    const float *row = p.x + ((size_t)b * p.C + c) * P;
    float *out = p.y + ((size_t)b * p.C + c) * P;
    for (int i = 0; i < P; ++i) {
        const float v = fmaf(__ldg(row + i), scale, shift);
        out[i] = p.relu ? fmaxf(v, 0.f) : v;
    }
}

__global__ void __launch_bounds__(NT) bn_bwd_reduce_small_kernel(const float *__restrict__ x,
                                                                 const float *__restrict__ dy,
                                                                 const float *__restrict__ y,
                                                                 const float *__restrict__ save_mean,
                                                                 const float *__restrict__ save_rstd,
                                                                 double *__restrict__ stats2, int C, int P) {
    const int c = blockIdx.x * NT + threadIdx.x, b = blockIdx.y;
    if (c >= C) return;
    const size_t off = ((size_t)b * C + c) * P;
    const float mean = save_mean[c], rstd = save_rstd[c];
    float s = 0.f, q = 0.f;
    for (int i = 0; i < P; ++i) {
        float g = __ldg(dy + off + i);
        if (y && !(__ldg(y + off + i) > 0.f)) g = 0.f;
        s += g;
        q += g * (__ldg(x + off + i) - mean);
    }
    q *= rstd;
    atomicAdd(stats2 + 2 * c, (double)s);
    atomicAdd(stats2 + 2 * c + 1, (double)q);
}

__global__ void __launch_bounds__(NT) bn_bwd_apply_small_kernel(const float *__restrict__ x,
                                                                const float *__restrict__ dy,
                                                                const float *__restrict__ y,
                                                                const float *__restrict__ save_mean,
                                                                const float *__restrict__ save_rstd,
                                                                const float *__restrict__ gamma,
                                                                const double *__restrict__ stats2, double inv_count,
                                                                float *__restrict__ dx, int C, int P) {
    const int c = blockIdx.x * NT + threadIdx.x, b = blockIdx.y;
    if (c >= C) return;
    const size_t off = ((size_t)b * C + c) * P;
    const float mean = save_mean[c], rstd = save_rstd[c];
    const float k0 = gamma[c] * rstd;
    const float m1 = (float)(stats2[2 * c] * inv_count);
    const float m2 = (float)(stats2[2 * c + 1] * inv_count) * rstd;
    for (int i = 0; i < P; ++i) {
        float g = __ldg(dy + off + i);
        if (y && !(__ldg(y + off + i) > 0.f)) g = 0.f;
        dx[off + i] = k0 * (g - m1 - (__ldg(x + off + i) - mean) * m2);
    }
}

// ---- BatchNorm -> max over K: statistics + per-(b,c,m) extrema in one read pass ------------------------------------
// y (B, C, M, K): thread <-> one (b,c,m) row of K contiguous values.
// vmax/vmin (B,C,M) f32; arg (B,C,M) int32 = argmax | (argmin << 16), first index on ties; stats[c] += (sum, sumsq).
__global__ void __launch_bounds__(NT) kmax_stats_kernel(const float *__restrict__ y, float *__restrict__ vmax,
                                                        float *__restrict__ vmin, int32_t *__restrict__ arg,
                                                        double *__restrict__ stats, int C, int M, int K) {
    const int c = blockIdx.y, b = blockIdx.z;
    const int m = blockIdx.x * NT + threadIdx.x;
    float s = 0.f, q = 0.f;
    if (m < M) {
        const size_t rowi = ((size_t)b * C + c) * M + m;
        const float *row = y + rowi * K;
        float mx = -INFINITY, mn = INFINITY;
        int ax = 0, an = 0;
        if ((K & 3) == 0 && (reinterpret_cast<uintptr_t>(row) & 15) == 0) {
            for (int k = 0; k < K; k += 4) {
                // lanes walk 32 different rows (stride K floats): keep the lines in L1 so the second half of every
                // 32-byte sector is not fetched again (the no-allocate streaming load would re-read it from L2)
                const float4 v = __ldg(reinterpret_cast<const float4 *>(row + k));
                const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    s += e[j];
                    q += e[j] * e[j];
                    if (e[j] > mx) { mx = e[j]; ax = k + j; }
                    if (e[j] < mn) { mn = e[j]; an = k + j; }
                }
            }
        } else {
            for (int k = 0; k < K; ++k) {
                const float e = __ldg(row + k);
                s += e;
                q += e * e;
                if (e > mx) { mx = e; ax = k; }
                if (e < mn) { mn = e; an = k; }
            }
        }
        vmax[rowi] = mx;
        vmin[rowi] = mn;
        arg[rowi] = ax | (an << 16);
    }
    if (stats) {
        block_sum2(s, q);
        if (threadIdx.x == 0) {
            atomicAdd(stats + 2 * c, (double)s);
            atomicAdd(stats + 2 * c + 1, (double)q);
        }
    }
}

// dy[b,c,m,k] = gsel[b,c,m] * [k == arg[b,c,m]] + a[c] + bcoef[c] * y[b,c,m,k]
__global__ void __launch_bounds__(NT) bnmax_bwd_kernel(const float *__restrict__ y, const float *__restrict__ gsel,
                                                       const int32_t *__restrict__ arg, const float *__restrict__ a,
                                                       const float *__restrict__ bcoef, float *__restrict__ dy, int C,
                                                       int M, int K) {
    const int c = blockIdx.y, b = blockIdx.z;
    const size_t row0 = ((size_t)b * C + c) * M;
    const float *src = y ? y + row0 * K : nullptr;
    float *out = dy + row0 * K;
    const float a0 = a ? a[c] : 0.f, b0 = bcoef ? bcoef[c] : 0.f;
    const long long total = (long long)M * K;
    if ((K & 3) == 0 && ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(out)) & 15) == 0) {
        // consecutive lanes own consecutive float4s of the (M,K) slab: fully coalesced reads and writes; the K/4 lanes of
        // one row read the same gsel / arg word (broadcast)
        for (long long e = ((long long)blockIdx.x * NT + threadIdx.x) * 4; e < total; e += (long long)gridDim.x * NT * 4) {
            const int m = (int)(e / K), k = (int)(e - (long long)m * K);
            const float g = __ldg(gsel + row0 + m);
            const int sel = __ldg(arg + row0 + m) - k;       // slot of the selected element relative to this float4
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (bcoef) v = apb_ld_nc_f4(src + e);
            v.x = fmaf(b0, v.x, a0) + (sel == 0 ? g : 0.f);
            v.y = fmaf(b0, v.y, a0) + (sel == 1 ? g : 0.f);
            v.z = fmaf(b0, v.z, a0) + (sel == 2 ? g : 0.f);
            v.w = fmaf(b0, v.w, a0) + (sel == 3 ? g : 0.f);
            *reinterpret_cast<float4 *>(out + e) = v;
        }
    } else {
        for (long long e = (long long)blockIdx.x * NT + threadIdx.x; e < total; e += (long long)gridDim.x * NT) {
            const int m = (int)(e / K), k = (int)(e - (long long)m * K);
            out[e] = fmaf(b0, bcoef ? __ldg(src + e) : 0.f, a0) + (__ldg(arg + row0 + m) == k ? __ldg(gsel + row0 + m) : 0.f);
        }
    }
}

// eval-mode BatchNorm (running statistics) is a per-channel affine map: y = act(scale[c] * x + shift[c]);
// with dy given instead (bwd = 1): dx = scale[c] * dy * [y_relu > 0]
__global__ void __launch_bounds__(NT) affine_kernel(const float *__restrict__ x, const float *__restrict__ y_relu,
                                                    const float *__restrict__ scale, const float *__restrict__ shift,
                                                    int relu, float *__restrict__ out, int C, long long P) {
    const int c = blockIdx.y, b = blockIdx.z;
    const size_t off = ((size_t)b * C + c) * P;
    const float sc = scale[c], sh = shift ? shift[c] : 0.f;
    const long long p0 = (long long)blockIdx.x * CHUNK;
    const long long p1 = min(P, p0 + CHUNK);
    const bool vec = (P & 3) == 0 && (((reinterpret_cast<uintptr_t>(x + off) | reinterpret_cast<uintptr_t>(out + off) |
                                        (y_relu ? reinterpret_cast<uintptr_t>(y_relu + off) : 0)) & 15) == 0);
    if (vec) {
        for (long long i = p0 + threadIdx.x * 4; i < p1; i += NT * 4) {
            float4 v = apb_ld_nc_f4(x + off + i);
            v.x = fmaf(v.x, sc, sh); v.y = fmaf(v.y, sc, sh); v.z = fmaf(v.z, sc, sh); v.w = fmaf(v.w, sc, sh);
            if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
            if (y_relu) {
                const float4 m = apb_ld_nc_f4(y_relu + off + i);
                v.x = m.x > 0.f ? v.x : 0.f; v.y = m.y > 0.f ? v.y : 0.f;
                v.z = m.z > 0.f ? v.z : 0.f; v.w = m.w > 0.f ? v.w : 0.f;
            }
            *reinterpret_cast<float4 *>(out + off + i) = v;
        }
    } else {
        for (long long i = p0 + threadIdx.x; i < p1; i += NT) {
            float v = fmaf(__ldg(x + off + i), sc, sh);
            if (relu) v = fmaxf(v, 0.f);
            if (y_relu && !(__ldg(y_relu + off + i) > 0.f)) v = 0.f;
            out[off + i] = v;
        }
    }
}

// short rows (P <= kSmallP): thread <-> (b, c) row, see bn_stats_small_kernel
__global__ void __launch_bounds__(NT) affine_small_kernel(const float *__restrict__ x, const float *__restrict__ y_relu,
                                                          const float *__restrict__ scale,
                                                          const float *__restrict__ shift, int relu,
                                                          float *__restrict__ out, int C, int P) {
    const int c = blockIdx.x * NT + threadIdx.x, b = blockIdx.y;
    if (c >= C) return;
    const size_t off = ((size_t)b * C + c) * P;
    const float sc = scale[c], sh = shift ? shift[c] : 0.f;
    for (int i = 0; i < P; ++i) {
        float v = fmaf(__ldg(x + off + i), sc, sh);
        if (relu) v = fmaxf(v, 0.f);
        if (y_relu && !(__ldg(y_relu + off + i) > 0.f)) v = 0.f;
        out[off + i] = v;
    }
}

bool grid_ok(int B, int C) { return B <= 65535 && C <= 65535; }


// ---- the small-tensor arithmetic around BatchNorm -> max over K (norm.BNMaxFn, fused_train.SaOneLayerFn) ---------------
// From the (C,2) sums to (mean, rstd, scale, shift), the running-statistics update, the max/min selection by the sign of
// the scale, the affine map and the ReLU: ONE kernel instead of ~28 element-wise torch launches per block and pass.
struct BnMaxFinParams {
    const double *stats;   // (C,2) sum, sum of squares over all ranks; nullptr: eval mode (running statistics)
    double inv_count, unbias;
    float eps, momentum;
    const float *gamma, *beta;
    float *running_mean, *running_var;   // read in eval mode, updated in training mode (may be nullptr there)
    const float *vmax, *vmin;
    const int32_t *arg;                  // kmax | kmin << 16
    int relu;
    float *out, *sel;
    int32_t *ksel;
    float *scale, *mean, *rstd;          // (C) saved for the backward pass
    int C, M;
};

__global__ void __launch_bounds__(NT) bnmax_finalize_kernel(const BnMaxFinParams p) {
    const int c = blockIdx.y, b = blockIdx.z;
    float mean, rstd;
    if (p.stats != nullptr) {
        const double mean_d = p.stats[2 * c] * p.inv_count;
        double var_d = p.stats[2 * c + 1] * p.inv_count - mean_d * mean_d;
        if (var_d < 0.0) var_d = 0.0;
        mean = (float)mean_d;
        rstd = (float)(1.0 / sqrt(var_d + (double)p.eps));
        if (b == 0 && blockIdx.x == 0 && threadIdx.x == 0 && p.running_mean != nullptr) {
            p.running_mean[c] = (1.f - p.momentum) * p.running_mean[c] + p.momentum * mean;
            p.running_var[c] = (1.f - p.momentum) * p.running_var[c] + p.momentum * (float)(var_d * p.unbias);
        }
    } else {
        mean = p.running_mean[c];
        rstd = rsqrtf(p.running_var[c] + p.eps);
    }
    const float scale = p.gamma[c] * rstd;
    const float shift = p.beta[c] - mean * scale;
    if (b == 0 && blockIdx.x == 0 && threadIdx.x == 0) {
        p.scale[c] = scale;
        p.mean[c] = mean;
        p.rstd[c] = rstd;
    }
    const bool pos = scale >= 0.f;
    const size_t base = ((size_t)b * p.C + c) * p.M;
    for (int m = blockIdx.x * NT + threadIdx.x; m < p.M; m += gridDim.x * NT) {
        const float v = pos ? p.vmax[base + m] : p.vmin[base + m];
        const int a = p.arg[base + m];
        float o = fmaf(v, scale, shift);
        if (p.relu) o = fmaxf(o, 0.f);
        p.out[base + m] = o;
        p.sel[base + m] = v;
        p.ksel[base + m] = pos ? (a & 0xFFFF) : (a >> 16);
    }
}

// gs = g * [out > 0] * scale ; sums (C,2) f64 += [sum g*mask, sum g*mask*xhat(sel)]
__global__ void __launch_bounds__(NT) bnmax_bwd_reduce_kernel(const float *__restrict__ g, const float *__restrict__ out,
                                                              const float *__restrict__ sel,
                                                              const float *__restrict__ scale,
                                                              const float *__restrict__ mean,
                                                              const float *__restrict__ rstd, float *__restrict__ gs,
                                                              double *__restrict__ sums, int C, int M) {
    const int c = blockIdx.y, b = blockIdx.z;
    const float sc = scale[c], mu = mean[c], rs = rstd[c];
    const size_t base = ((size_t)b * C + c) * M;
    float s = 0.f, q = 0.f;
    for (int m = blockIdx.x * NT + threadIdx.x; m < M; m += gridDim.x * NT) {
        float gv = g[base + m];
        if (out != nullptr && !(out[base + m] > 0.f)) gv = 0.f;
        gs[base + m] = gv * sc;
        s += gv;
        q = fmaf(gv, (sel[base + m] - mu) * rs, q);
    }
    block_sum2(s, q);
    if (threadIdx.x == 0) {
        atomicAdd(sums + 2 * c, (double)s);
        atomicAdd(sums + 2 * c + 1, (double)q);
    }
}

// per channel: BatchNorm-backward coefficients of the dense terms (training) and the parameter gradients
__global__ void bnmax_bwd_coef_kernel(const double *__restrict__ tot, const double *__restrict__ local,
                                      const float *__restrict__ scale, const float *__restrict__ mean,
                                      const float *__restrict__ rstd, double inv_count, float *__restrict__ a,
                                      float *__restrict__ bcoef, float *__restrict__ dgamma, float *__restrict__ dbeta,
                                      int C) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    dbeta[c] = (float)local[2 * c];
    dgamma[c] = (float)local[2 * c + 1];
    if (a != nullptr) {
        const float bc = (float)(-(double)scale[c] * (double)rstd[c] * tot[2 * c + 1] * inv_count);
        bcoef[c] = bc;
        a[c] = (float)(-(double)scale[c] * tot[2 * c] * inv_count) - bc * mean[c];
    }
}

}  // namespace

// vmax[r] = max_p y[r,p], arg[r] = first p attaining it.   One warp per row, 16-byte loads.
// Used for "1x1 conv -> ReLU -> max over the points" heads (point_discriminator.py:166-191): since ReLU is monotone,
// max(relu(y)) = relu(max(y)), so the ReLU'd activation is never materialised and the backward pass only touches the
// arg-max columns.
__global__ void __launch_bounds__(256) rowmax_kernel(const float *__restrict__ y, float *__restrict__ vmax,
                                                      int32_t *__restrict__ arg, long long rows, int P) {
    const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int lane = threadIdx.x & 31;
    const float *row = y + r * P;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    if ((P & 3) == 0 && ((reinterpret_cast<uintptr_t>(row) & 15) == 0)) {
        for (int p = lane * 4; p < P; p += 128) {
            const float4 v = apb_ld_nc_f4(row + p);
            if (v.x > best) { best = v.x; bi = p; }
            if (v.y > best) { best = v.y; bi = p + 1; }
            if (v.z > best) { best = v.z; bi = p + 2; }
            if (v.w > best) { best = v.w; bi = p + 3; }
        }
    } else {
        for (int p = lane; p < P; p += 32) {
            const float v = __ldg(row + p);
            if (v > best) { best = v; bi = p; }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (lane == 0) {
        vmax[r] = best;
        arg[r] = bi == 0x7fffffff ? 0 : bi;
    }
}


extern "C" int apb_bn_stats_f32(const float *x, double *stats, int B, int C, long long P, apb_stream_t stream) {
    if (B < 0 || C < 0 || P < 0) return APB_E_INVALID;
    if (B == 0 || C == 0 || P == 0) return APB_OK;
    if (!x || !stats) return APB_E_INVALID;
    if (!grid_ok(B, C)) return APB_E_UNSUPPORTED;
    if (P <= kSmallP) {
        bn_stats_small_kernel<<<dim3(apb_div_up(C, NT), B), NT, 0, (cudaStream_t)stream>>>(x, stats, C, (int)P);
        APB_CHECK_LAUNCH();
        return APB_OK;
    }
    dim3 grid(apb_div_up(P, CHUNK), C, B);
    bn_stats_kernel<<<grid, NT, 0, (cudaStream_t)stream>>>(x, stats, C, P);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_bn_apply_f32(const float *x, const double *stats, const float *gamma, const float *beta, float eps,
                                double count, int relu, float *y, float *save_mean, float *save_rstd,
                                float *running_mean, float *running_var, float momentum, int B, int C, long long P,
                                apb_stream_t stream) {
    if (B < 0 || C < 0 || P < 0 || count < 1.0) return APB_E_INVALID;
    if (B == 0 || C == 0 || P == 0) return APB_OK;
    if (!x || !stats || !gamma || !beta || !y || !save_mean || !save_rstd) return APB_E_INVALID;
    if ((running_mean == nullptr) != (running_var == nullptr)) return APB_E_INVALID;
    if (!grid_ok(B, C)) return APB_E_UNSUPPORTED;
    BnApplyParams p{x, y, stats, gamma, beta, eps, 1.0 / count, count > 1.0 ? count / (count - 1.0) : 1.0, relu,
                    save_mean, save_rstd, running_mean, running_var, momentum, C, P};
    if (P <= kSmallP) {
        bn_apply_small_kernel<<<dim3(apb_div_up(C, NT), B), NT, 0, (cudaStream_t)stream>>>(p);
        APB_CHECK_LAUNCH();
        return APB_OK;
    }
    dim3 grid(apb_div_up(P, CHUNK), C, B);
    bn_apply_kernel<<<grid, NT, 0, (cudaStream_t)stream>>>(p);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_bn_bwd_reduce_f32(const float *x, const float *dy, const float *y_relu, const float *save_mean,
                                     const float *save_rstd, double *stats2, int B, int C, long long P,
                                     apb_stream_t stream) {
    if (B < 0 || C < 0 || P < 0) return APB_E_INVALID;
    if (B == 0 || C == 0 || P == 0) return APB_OK;
    if (!x || !dy || !save_mean || !save_rstd || !stats2) return APB_E_INVALID;
    if (!grid_ok(B, C)) return APB_E_UNSUPPORTED;
    if (P <= kSmallP) {
        bn_bwd_reduce_small_kernel<<<dim3(apb_div_up(C, NT), B), NT, 0, (cudaStream_t)stream>>>(
            x, dy, y_relu, save_mean, save_rstd, stats2, C, (int)P);
        APB_CHECK_LAUNCH();
        return APB_OK;
    }
    dim3 grid(apb_div_up(P, CHUNK), C, B);
    bn_bwd_reduce_kernel<<<grid, NT, 0, (cudaStream_t)stream>>>(x, dy, y_relu, save_mean, save_rstd, stats2, C, P);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_bn_bwd_apply_f32(const float *x, const float *dy, const float *y_relu, const float *save_mean,
                                    const float *save_rstd, const float *gamma, const double *stats2, double count,
                                    float *dx, int B, int C, long long P, apb_stream_t stream) {
    if (B < 0 || C < 0 || P < 0 || count < 1.0) return APB_E_INVALID;
    if (B == 0 || C == 0 || P == 0) return APB_OK;
    if (!x || !dy || !save_mean || !save_rstd || !gamma || !stats2 || !dx) return APB_E_INVALID;
    if (!grid_ok(B, C)) return APB_E_UNSUPPORTED;
    if (P <= kSmallP) {
        bn_bwd_apply_small_kernel<<<dim3(apb_div_up(C, NT), B), NT, 0, (cudaStream_t)stream>>>(
            x, dy, y_relu, save_mean, save_rstd, gamma, stats2, 1.0 / count, dx, C, (int)P);
        APB_CHECK_LAUNCH();
        return APB_OK;
    }
    dim3 grid(apb_div_up(P, CHUNK), C, B);
    bn_bwd_apply_kernel<<<grid, NT, 0, (cudaStream_t)stream>>>(x, dy, y_relu, save_mean, save_rstd, gamma, stats2,
                                                               1.0 / count, dx, C, P);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_kmax_stats_f32(const float *y, float *vmax, float *vmin, int32_t *arg, double *stats, int B, int C,
                                  int M, int K, apb_stream_t stream) {
    if (B < 0 || C < 0 || M < 0 || K <= 0 || K > 32767) return APB_E_INVALID;
    if (B == 0 || C == 0 || M == 0) return APB_OK;
    if (!y || !vmax || !vmin || !arg) return APB_E_INVALID;
    if (!grid_ok(B, C)) return APB_E_UNSUPPORTED;
    dim3 grid(apb_div_up(M, NT), C, B);
    kmax_stats_kernel<<<grid, NT, 0, (cudaStream_t)stream>>>(y, vmax, vmin, arg, stats, C, M, K);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_bnmax_bwd_f32(const float *y, const float *gsel, const int32_t *arg, const float *a,
                                 const float *bcoef, float *dy, int B, int C, int M, int K, apb_stream_t stream) {
    if (B < 0 || C < 0 || M < 0 || K <= 0) return APB_E_INVALID;
    if (B == 0 || C == 0 || M == 0) return APB_OK;
    if (!gsel || !arg || !dy || (bcoef && !y)) return APB_E_INVALID;
    if (!grid_ok(B, C)) return APB_E_UNSUPPORTED;
    dim3 grid(min(apb_div_up((long long)M * K, NT * 4), 64), C, B);
    bnmax_bwd_kernel<<<grid, NT, 0, (cudaStream_t)stream>>>(y, gsel, arg, a, bcoef, dy, C, M, K);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_affine_f32(const float *x, const float *mask_relu, const float *scale, const float *shift, int relu,
                              float *out, int B, int C, long long P, apb_stream_t stream) {
    if (B < 0 || C < 0 || P < 0) return APB_E_INVALID;
    if (B == 0 || C == 0 || P == 0) return APB_OK;
    if (!x || !scale || !out) return APB_E_INVALID;
    if (!grid_ok(B, C)) return APB_E_UNSUPPORTED;
    if (P <= kSmallP) {
        affine_small_kernel<<<dim3(apb_div_up(C, NT), B), NT, 0, (cudaStream_t)stream>>>(x, mask_relu, scale, shift, relu,
                                                                                         out, C, (int)P);
        APB_CHECK_LAUNCH();
        return APB_OK;
    }
    dim3 grid(apb_div_up(P, CHUNK), C, B);
    affine_kernel<<<grid, NT, 0, (cudaStream_t)stream>>>(x, mask_relu, scale, shift, relu, out, C, P);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_rowmax_f32(const float *y, float *vmax, int32_t *arg, long long rows, int P, apb_stream_t stream) {
    if (rows < 0 || P < 0) return APB_E_INVALID;
    if (rows == 0) return APB_OK;
    if (!y || !vmax || !arg || P == 0) return APB_E_INVALID;
    if (rows > 8LL * 2147483647LL) return APB_E_UNSUPPORTED;
    rowmax_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, (cudaStream_t)stream>>>(y, vmax, arg, rows, P);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_bnmax_finalize_f32(const double *stats, double count, float eps, float momentum, const float *gamma,
                                      const float *beta, float *running_mean, float *running_var, const float *vmax,
                                      const float *vmin, const int32_t *arg, int relu, float *out, float *sel,
                                      int32_t *ksel, float *scale, float *mean, float *rstd, int B, int C, int M,
                                      apb_stream_t stream) {
    if (B < 0 || C <= 0 || M < 0) return APB_E_INVALID;
    if (B == 0 || M == 0) return APB_OK;
    if (!gamma || !beta || !vmax || !vmin || !arg || !out || !sel || !ksel || !scale || !mean || !rstd) return APB_E_INVALID;
    if (stats == nullptr && (!running_mean || !running_var)) return APB_E_INVALID;
    if (B > 65535 || C > 65535) return APB_E_UNSUPPORTED;
    BnMaxFinParams p{};
    p.stats = stats;
    p.inv_count = count > 0 ? 1.0 / count : 0.0;
    p.unbias = count > 1 ? count / (count - 1.0) : 1.0;
    p.eps = eps; p.momentum = momentum; p.gamma = gamma; p.beta = beta;
    p.running_mean = running_mean; p.running_var = running_var;
    p.vmax = vmax; p.vmin = vmin; p.arg = arg; p.relu = relu; p.out = out; p.sel = sel; p.ksel = ksel;
    p.scale = scale; p.mean = mean; p.rstd = rstd; p.C = C; p.M = M;
    int gx = apb_div_up(M, NT);
    if (gx > 8) gx = 8;
    bnmax_finalize_kernel<<<dim3(gx, C, B), NT, 0, (cudaStream_t)stream>>>(p);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_bnmax_bwd_reduce_f32(const float *g, const float *out_relu, const float *sel, const float *scale,
                                        const float *mean, const float *rstd, float *gs, double *sums, int B, int C, int M,
                                        apb_stream_t stream) {
    if (B < 0 || C <= 0 || M < 0) return APB_E_INVALID;
    if (B == 0 || M == 0) return APB_OK;
    if (!g || !sel || !scale || !mean || !rstd || !gs || !sums) return APB_E_INVALID;
    if (B > 65535 || C > 65535) return APB_E_UNSUPPORTED;
    int gx = apb_div_up(M, NT * 4);
    if (gx > 8) gx = 8;
    bnmax_bwd_reduce_kernel<<<dim3(gx, C, B), NT, 0, (cudaStream_t)stream>>>(g, out_relu, sel, scale, mean, rstd, gs, sums,
                                                                             C, M);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_bnmax_bwd_coef_f32(const double *tot, const double *local, const float *scale, const float *mean,
                                      const float *rstd, double count, float *a, float *bcoef, float *dgamma,
                                      float *dbeta, int C, apb_stream_t stream) {
    if (C <= 0) return APB_E_INVALID;
    if (!tot || !local || !scale || !mean || !rstd || !dgamma || !dbeta || ((a == nullptr) != (bcoef == nullptr)))
        return APB_E_INVALID;
    bnmax_bwd_coef_kernel<<<apb_div_up(C, 128), 128, 0, (cudaStream_t)stream>>>(tot, local, scale, mean, rstd,
                                                                                count > 0 ? 1.0 / count : 0.0, a, bcoef,
                                                                                dgamma, dbeta, C);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

// ---- eval-mode fold and small conversions: one launch instead of 4-5 element-wise ones per BatchNorm layer ------------
__global__ void bn_fold_kernel(const float *__restrict__ gamma, const float *__restrict__ beta,
                               const float *__restrict__ mean, const float *__restrict__ var, float eps,
                               float *__restrict__ scale, float *__restrict__ shift, float *__restrict__ rstd, int C) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const float r = rsqrtf(var[c] + eps);
    const float s = gamma[c] * r;
    rstd[c] = r;
    scale[c] = s;
    shift[c] = beta[c] - mean[c] * s;
}

// (C,2) sums -> mean, rstd, scale = gamma*rstd, shift = beta - mean*scale, running-statistics update (training-mode BatchNorm
// whose normalised output is never materialised: the first BatchNorm of the fused two-conv set abstraction, sa_train2.cu)
__global__ void bn_finalize_kernel(const double *__restrict__ stats, double inv_count, double unbias, float eps,
                                   float momentum, const float *__restrict__ gamma, const float *__restrict__ beta,
                                   float *__restrict__ running_mean, float *__restrict__ running_var,
                                   float *__restrict__ scale, float *__restrict__ shift, float *__restrict__ mean_out,
                                   float *__restrict__ rstd_out, int C) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    const double mean_d = stats[2 * c] * inv_count;
    double var_d = stats[2 * c + 1] * inv_count - mean_d * mean_d;
    if (var_d < 0.0) var_d = 0.0;
    const float mean = (float)mean_d;
    const float rstd = (float)(1.0 / sqrt(var_d + (double)eps));
    if (running_mean != nullptr) {
        running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * mean;
        running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)(var_d * unbias);
    }
    const float sc = gamma[c] * rstd;
    scale[c] = sc;
    shift[c] = beta[c] - mean * sc;
    mean_out[c] = mean;
    rstd_out[c] = rstd;
}

__global__ void sums_split_kernel(const double *__restrict__ sums, float *__restrict__ first, float *__restrict__ second,
                                  int C) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= C) return;
    first[c] = (float)sums[2 * c];
    second[c] = (float)sums[2 * c + 1];
}

extern "C" int apb_bn_fold_f32(const float *gamma, const float *beta, const float *running_mean, const float *running_var,
                               float eps, float *scale, float *shift, float *rstd, int C, apb_stream_t stream) {
    if (C <= 0) return APB_E_INVALID;
    if (!gamma || !beta || !running_mean || !running_var || !scale || !shift || !rstd) return APB_E_INVALID;
    bn_fold_kernel<<<apb_div_up(C, 128), 128, 0, (cudaStream_t)stream>>>(gamma, beta, running_mean, running_var, eps, scale,
                                                                         shift, rstd, C);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_bn_finalize_f32(const double *stats, double count, float eps, float momentum, const float *gamma,
                                   const float *beta, float *running_mean, float *running_var, float *scale, float *shift,
                                   float *mean, float *rstd, int C, apb_stream_t stream) {
    if (C <= 0 || count <= 0) return APB_E_INVALID;
    if (!stats || !gamma || !beta || !scale || !shift || !mean || !rstd || ((running_mean == nullptr) != (running_var == nullptr)))
        return APB_E_INVALID;
    bn_finalize_kernel<<<apb_div_up(C, 128), 128, 0, (cudaStream_t)stream>>>(
        stats, 1.0 / count, count > 1 ? count / (count - 1) : 1.0, eps, momentum, gamma, beta, running_mean, running_var, scale,
        shift, mean, rstd, C);
    APB_CHECK_LAUNCH();
    return APB_OK;
}

extern "C" int apb_sums_split_f32(const double *sums, float *first, float *second, int C, apb_stream_t stream) {
    if (C <= 0) return APB_E_INVALID;
    if (!sums || !first || !second) return APB_E_INVALID;
    sums_split_kernel<<<apb_div_up(C, 128), 128, 0, (cudaStream_t)stream>>>(sums, first, second, C);
    APB_CHECK_LAUNCH();
    return APB_OK;
}
