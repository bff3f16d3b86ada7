// Per-step operators around the right-hand side -- sm_100a kernels, FP64, HBM-bound.
//
//   wxb_expfilter       per-step exponential modal filter  ((sqrtG Q) @ expfilter_new) * inv_sqrtG
//                       wx_factory/geometry/operators.py:101-119,257-261 (applied by simulation.py:148), as three
//                       sum-factorised 1-D passes in shared memory instead of the dense (P x P) GEMM
//   wxb_vec_lincomb3    ((a x) + (b y)) + ((c z) dt): the stage combinations of SSP-RK3
//                       (wx_factory/integrators/tvdrk3.py:13-17) and Ros2's b = A(Q) + rhs dt
//                       (wx_factory/integrators/ros2.py:30), rounded operation by operation like the numpy expressions
//   wxb_vec_nonfinite   the NaN check of simulation.py:399-408
//   wxb_vec_multidot    out[i][c] = V[i] . W[c]: the Gram-Schmidt products of KIOPS (wx_factory/solvers/kiops.py:190-197,
//                       201-205), PMEX (wx_factory/solvers/pmex.py:150-230) and of the 1-sync FGMRES
//                       (wx_factory/solvers/fgmres.py:38); fixed-order two-stage reduction (bit-reproducible run to run)
//   wxb_vec_combine     out = (b0 y + sum_i c[i] V[i]) * s: kiops.py:199,208,278,286 / Krylov solution updates
//   wxb_vec_reduce      sum |x| (kiops.py:111-116) and max |y/x| over x != 0 (fgmres.py:257-262)
//   wxb_vec_gs1sync     the row updates of fgmres.py:60-69 (_ortho_1_sync_igs) in one pass over the basis
//
// The vector kernels are free functions of the C ABI (no handle): Krylov vectors are plain device arrays.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

#include "wxb_internal.cuh"

namespace {

void set_err(wxb_handle* h, const char* fmt, ...) {  // same convention as wxb_api.cu: message kept on the handle
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(h->err, sizeof(h->err), fmt, ap);
    va_end(ap);
}

constexpr int VEC_THREADS = 256;
constexpr int VEC_BLOCKS = 148 * 4;  // persistent grid-stride blocks: a multiple of the SM count
constexpr int MD_ROWS = 8;           // basis rows per pass of the multi-dot kernel (register accumulators); measured on B200 at 53 M
                                     // values: 8 rows 2.3 ms for 22 x 2 products, 12 rows 4.9 ms (12 widely spaced streams per thread)

// ------------------------------------------------------------------------------------------------------
// exponential filter.  One thread owns one LINE of N nodes of an element in each of the three passes, so that every
// pass reads and writes each value once (a thread per node reads each value N times from shared memory, which is what
// bound the round-1 kernel: 120 B of shared-memory traffic per node and variable against 32 B here).
//   pass x: thread (j,k) loads its line of q and sqrtG straight from global memory (N contiguous doubles), filters it
//           in registers, stores it to shared memory;
//   pass y: thread (i,k) filters its line in shared memory, in place;
//   pass z: thread (i,j) reads its line, filters it, scales by 1/sqrtG and writes the result to global memory.
// Plane stride N*N + N: the 16 lines (i,k) of an element then fall in 16 different 8-byte banks in pass y (n = 4).
template <int N>
struct FiltCfg {
    static constexpr int P = N * N * N, NV = 5, L = N * N;           // L lines per element and pass
    static constexpr int EB = (256 / L) < 1 ? 1 : 256 / L;           // elements per block
    static constexpr int THREADS = EB * L;
    static constexpr int S = N * N + N;                              // padded k-plane stride
    static constexpr int ELEM = N * S;                               // doubles per element and variable
    static constexpr size_t SMEM = (size_t)EB * NV * ELEM * sizeof(double);
};

// the lines of an element exchange values only among themselves: when a warp holds whole elements (N = 2, 4) a warp
// barrier is enough and the warps of a block run the three passes independently of one another
template <int L>
__device__ __forceinline__ void filt_sync() {
    if constexpr (32 % L == 0) __syncwarp();
    else __syncthreads();
}

template <int N, bool VEC>  // VEC: N even and 16-byte aligned arrays, the pass-x loads are 128-bit
__global__ void __launch_bounds__(FiltCfg<N>::THREADS) expfilter_kernel(const EvalArgs a, const double* __restrict__ filt,
                                                                        const double* __restrict__ metric) {
    using C = FiltCfg<N>;
    constexpr int P = C::P, NV = C::NV, L = C::L, S = C::S;
    extern __shared__ __align__(16) double s_x[];  // [EB][NV][N planes][S]
    const int le = threadIdx.x / L, line = threadIdx.x % L;
    const int e = blockIdx.x * C::EB + le;
    const bool active = e < a.E;
    const int la = line % N, lb = line / N;  // the two fixed coordinates of this thread's lines
    double f[N][N];                          // the filter matrix, kept in registers by every thread
#pragma unroll
    for (int r = 0; r < N; ++r)
#pragma unroll
        for (int m = 0; m < N; ++m) f[r][m] = filt[r * N + m];
    double* se = s_x + (size_t)le * NV * C::ELEM;
    long long qo = 0, vst = 0;
    const double* mrow = nullptr;
    if (active) {
        const ElemPos ep = decode_elem(a.st, e, a.nz, a.nx, P, NV);
        qo = ep.qoff;
        vst = ep.vstride;
        mrow = metric + (long long)e * (35 * P) + 27 * P;  // row sqrtG of the element-major metric block
        // pass x: line (j = la, k = lb), nodes lb N^2 + la N + 0..N-1
        const int n0 = lb * N * N + la * N;
        double sg[N], x[NV][N];
        if constexpr (VEC) {
#pragma unroll
            for (int i = 0; i < N; i += 2) {
                const double2 t = __ldg(reinterpret_cast<const double2*>(mrow + n0 + i));
                sg[i] = t.x, sg[i + 1] = t.y;
            }
#pragma unroll
            for (int v = 0; v < NV; ++v)
#pragma unroll
                for (int i = 0; i < N; i += 2) {
                    const double2 t = *reinterpret_cast<const double2*>(a.q + qo + v * vst + n0 + i);
                    x[v][i] = t.x, x[v][i + 1] = t.y;
                }
        } else {
#pragma unroll
            for (int i = 0; i < N; ++i) sg[i] = mrow[n0 + i];
#pragma unroll
            for (int v = 0; v < NV; ++v)
#pragma unroll
                for (int i = 0; i < N; ++i) x[v][i] = a.q[qo + v * vst + n0 + i];
        }
#pragma unroll
        for (int v = 0; v < NV; ++v) {
#pragma unroll
            for (int i = 0; i < N; ++i) x[v][i] *= sg[i];
#pragma unroll
            for (int r = 0; r < N; ++r) {
                double t = 0.0;
#pragma unroll
                for (int m = 0; m < N; ++m) t = fma(f[r][m], x[v][m], t);
                se[v * C::ELEM + lb * S + la * N + r] = t;
            }
        }
    }
    filt_sync<L>();
    if (active) {
        // pass y: line (i = la, k = lb), in place
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            double* ln = se + v * C::ELEM + lb * S + la;
            double x[N];
#pragma unroll
            for (int m = 0; m < N; ++m) x[m] = ln[m * N];
#pragma unroll
            for (int r = 0; r < N; ++r) {
                double t = 0.0;
#pragma unroll
                for (int m = 0; m < N; ++m) t = fma(f[r][m], x[m], t);
                ln[r * N] = t;
            }
        }
    }
    filt_sync<L>();
    if (active) {
        // pass z: line (i = la, j = lb); consecutive threads write consecutive nodes of each k plane
        double inv[N];
#pragma unroll
        for (int k = 0; k < N; ++k) inv[k] = 1.0 / mrow[k * N * N + line];  // metric.inv_sqrtG_new (metric3d.py:1157)
#pragma unroll
        for (int v = 0; v < NV; ++v) {
            const double* ln = se + v * C::ELEM + line;
            double x[N];
#pragma unroll
            for (int m = 0; m < N; ++m) x[m] = ln[m * S];
#pragma unroll
            for (int r = 0; r < N; ++r) {
                double t = 0.0;
#pragma unroll
                for (int m = 0; m < N; ++m) t = fma(f[r][m], x[m], t);
                a.out[qo + v * vst + r * N * N + line] = t * inv[r];
            }
        }
    }
}

template <int N>
int launch_expfilter(const EvalArgs& a, const double* filt, const double* metric, cudaStream_t s) {
    using C = FiltCfg<N>;
    const bool vec = N % 2 == 0 && (((uintptr_t)a.q | (uintptr_t)metric) & 15) == 0;
    auto kern = vec ? expfilter_kernel<N, (N % 2 == 0)> : expfilter_kernel<N, false>;
    if (C::SMEM > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM) != cudaSuccess)
        return -1;
    kern<<<(a.E + C::EB - 1) / C::EB, C::THREADS, C::SMEM, s>>>(a, filt, metric);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// ------------------------------------------------------------------------------------------------------
// ((a x) + (b y)) + ((c z) dt), every operation rounded on its own (no fused multiply-add), so that the result is
// the bit pattern numpy produces for  a*x + b*y + c*z*dt  evaluated left to right.  y == nullptr: (a x) + ((c z) dt).
__global__ void __launch_bounds__(VEC_THREADS) lincomb3_kernel(long long n, double ca, const double* __restrict__ x, double cb,
                                                              const double* __restrict__ y, double cc,
                                                              const double* __restrict__ z, double dt, double* __restrict__ out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) {
        double s = ca == 1.0 ? x[i] : __dmul_rn(ca, x[i]);
        if (y) s = __dadd_rn(s, __dmul_rn(cb, y[i]));
        out[i] = __dadd_rn(s, __dmul_rn(__dmul_rn(cc, z[i]), dt));
    }
}

__global__ void __launch_bounds__(VEC_THREADS) nonfinite_kernel(long long n, const double* __restrict__ x, unsigned long long* count) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    unsigned int bad = 0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) bad += !isfinite(x[i]);
    bad = __reduce_add_sync(0xffffffffu, bad);
    if ((threadIdx.x & 31) == 0 && bad) atomicAdd(count, (unsigned long long)bad);  // integer count: order-independent
}

// ------------------------------------------------------------------------------------------------------
// multi-dot, stage 1: partial[block][i][c] = sum over the block's grid-stride entries of V[i][x] W[c][x]
// (m <= MD_ROWS rows per launch, K = 1 or 2 right-hand vectors); warp-shuffle + shared-memory reduction in a fixed order
// ROWS = rows per pass: 8, or 2 for the dots / norms / short incomplete orthogonalisations of KIOPS, where the entry loop
// is unrolled by 4 instead so that enough loads are in flight per thread.
template <int K, int VEC, int ROWS>
__global__ void __launch_bounds__(VEC_THREADS) multidot_partial(long long n, int m, const double* __restrict__ V, long long ldv,
                                                               const double* __restrict__ W, long long ldw, double* __restrict__ partial,
                                                               int mtot, int row0) {
    constexpr int MD_ROWS = ROWS;
    constexpr int UNR = ROWS <= 2 ? 4 : 1;
    double acc[MD_ROWS][K];
#pragma unroll
    for (int i = 0; i < MD_ROWS; ++i)
#pragma unroll
        for (int c = 0; c < K; ++c) acc[i][c] = 0.0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long gid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (VEC == 2) {  // rows 16-byte aligned: two entries per 128-bit load (half the load instructions per byte)
        const long long nv = n >> 1;
#pragma unroll UNR
        for (long long x = gid; x < nv; x += stride) {
            double2 w[K];
#pragma unroll
            for (int c = 0; c < K; ++c) w[c] = reinterpret_cast<const double2*>(W + c * ldw)[x];
#pragma unroll
            for (int i = 0; i < MD_ROWS; ++i) {
                if (i < m) {
                    const double2 v = reinterpret_cast<const double2*>(V + i * ldv)[x];
#pragma unroll
                    for (int c = 0; c < K; ++c) acc[i][c] = fma(v.y, w[c].y, fma(v.x, w[c].x, acc[i][c]));
                }
            }
        }
        if ((n & 1) && gid == 0) {  // odd length: the last entry
#pragma unroll
            for (int i = 0; i < MD_ROWS; ++i)
                if (i < m)
#pragma unroll
                    for (int c = 0; c < K; ++c) acc[i][c] = fma(V[i * ldv + n - 1], W[c * ldw + n - 1], acc[i][c]);
        }
    } else {
#pragma unroll UNR
        for (long long x = gid; x < n; x += stride) {
            double w[K];
#pragma unroll
            for (int c = 0; c < K; ++c) w[c] = W[c * ldw + x];
#pragma unroll
            for (int i = 0; i < MD_ROWS; ++i) {
                if (i < m) {
                    const double v = V[i * ldv + x];
#pragma unroll
                    for (int c = 0; c < K; ++c) acc[i][c] = fma(v, w[c], acc[i][c]);
                }
            }
        }
    }
    __shared__ double s_red[VEC_THREADS / 32][MD_ROWS][K];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < MD_ROWS; ++i)
#pragma unroll
        for (int c = 0; c < K; ++c) {
            double v = acc[i][c];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
            if (lane == 0) s_red[warp][i][c] = v;
        }
    __syncthreads();
    if (threadIdx.x < MD_ROWS * K) {
        const int i = threadIdx.x / K, c = threadIdx.x % K;
        if (i < m) {
            double v = 0.0;
#pragma unroll
            for (int wgt = 0; wgt < VEC_THREADS / 32; ++wgt) v += s_red[wgt][i][c];
            partial[((long long)blockIdx.x * mtot + row0 + i) * K + c] = v;
        }
    }
}

// multi-dot, stage 1, ALL rows of a launch in one pass over W: the thread keeps MR x K accumulators and walks the rows in
// groups of 8 (at most 8 row streams in flight per thread: 12 measured 2x slower), so the right-hand vectors are read once
// per launch instead of once per group of 8 rows (22 x 2 products: 24 row reads instead of 28).  128-bit loads only
// (rows 16-byte aligned); same per-thread summation order as multidot_partial, same block reduction.
template <int K, int MR>
__global__ void __launch_bounds__(VEC_THREADS, MR <= 16 ? 2 : 1) multidot_allrows(long long n, int m, const double* __restrict__ V, long long ldv,
                                                                                  const double* __restrict__ W, long long ldw,
                                                                                  double* __restrict__ partial, int mtot, int row0) {
    double acc[MR][K];
#pragma unroll
    for (int i = 0; i < MR; ++i)
#pragma unroll
        for (int c = 0; c < K; ++c) acc[i][c] = 0.0;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long gid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long nv = n >> 1;
    for (long long x = gid; x < nv; x += stride) {
        double2 w[K];
#pragma unroll
        for (int c = 0; c < K; ++c) w[c] = reinterpret_cast<const double2*>(W + c * ldw)[x];
#pragma unroll
        for (int g = 0; g < MR; g += 8) {
            double2 v[8];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (g + i < MR && g + i < m) v[i] = reinterpret_cast<const double2*>(V + (g + i) * ldv)[x];
#pragma unroll
            for (int i = 0; i < 8; ++i)
                if (g + i < MR && g + i < m) {
#pragma unroll
                    for (int c = 0; c < K; ++c) acc[g + i][c] = fma(v[i].y, w[c].y, fma(v[i].x, w[c].x, acc[g + i][c]));
                }
        }
    }
    if ((n & 1) && gid == 0) {  // odd length: the last entry
#pragma unroll
        for (int i = 0; i < MR; ++i)
            if (i < m)
#pragma unroll
                for (int c = 0; c < K; ++c) acc[i][c] = fma(V[i * ldv + n - 1], W[c * ldw + n - 1], acc[i][c]);
    }
    __shared__ double s_red[VEC_THREADS / 32][MR][K];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < MR; ++i)
#pragma unroll
        for (int c = 0; c < K; ++c) {
            double v = acc[i][c];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
            if (lane == 0) s_red[warp][i][c] = v;
        }
    __syncthreads();
    if (threadIdx.x < MR * K) {
        const int i = threadIdx.x / K, c = threadIdx.x % K;
        if (i < m) {
            double v = 0.0;
#pragma unroll
            for (int wgt = 0; wgt < VEC_THREADS / 32; ++wgt) v += s_red[wgt][i][c];
            partial[((long long)blockIdx.x * mtot + row0 + i) * K + c] = v;
        }
    }
}

// stage 2: out[i][c] = sum over blocks, in block order
__global__ void multidot_final(int nblocks, int mk, const double* __restrict__ partial, double* __restrict__ out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= mk) return;
    double v = 0.0;
    for (int b = 0; b < nblocks; ++b) v += partial[(long long)b * mk + t];
    out[t] = v;
}

// scalar reductions the solvers need once per call / restart cycle, two stages in a fixed order like the multi-dot:
//   kind 0: sum |x|                       (1-norms of the phi-function vectors, kiops.py:111-116, pmex.py:91-98)
//   kind 1: max |y / x| over x != 0, -1 when x has no non-zero entry   (stagnation test, fgmres.py:257-262)
template <int KIND>
__global__ void __launch_bounds__(VEC_THREADS) reduce_partial(long long n, const double* __restrict__ x, const double* __restrict__ y,
                                                             double* __restrict__ partial) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    double acc = KIND == 0 ? 0.0 : -1.0;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += stride) {
        if (KIND == 0) acc += fabs(x[i]);
        else if (x[i] != 0.0) acc = fmax(acc, fabs(y[i] / x[i]));
    }
    __shared__ double s_red[VEC_THREADS / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double t = __shfl_down_sync(0xffffffffu, acc, o);
        acc = KIND == 0 ? acc + t : fmax(acc, t);
    }
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double v = s_red[0];
        for (int w = 1; w < VEC_THREADS / 32; ++w) v = KIND == 0 ? v + s_red[w] : fmax(v, s_red[w]);
        partial[blockIdx.x] = v;
    }
}

template <int KIND>
__global__ void reduce_final(int nblocks, const double* __restrict__ partial, double* __restrict__ out) {
    if (blockIdx.x || threadIdx.x) return;
    double v = partial[0];
    for (int b = 1; b < nblocks; ++b) v = KIND == 0 ? v + partial[b] : fmax(v, partial[b]);
    *out = v;
}

struct Coefs {
    double c[WXB_VEC_MAX_ROWS];
};

// out = (b0 y + sum_i c[i] V[i]) * s   (y may be null; out may alias y or a row of V that is not summed over)
__global__ void __launch_bounds__(VEC_THREADS) combine_kernel(long long n, int m, const Coefs cf, const double* __restrict__ V, long long ldv,
                                                             double b0, const double* y, double s, double* out) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long x = blockIdx.x * (long long)blockDim.x + threadIdx.x; x < n; x += stride) {
        double acc = 0.0;
#pragma unroll 4
        for (int i = 0; i < m; ++i) acc = fma(cf.c[i], V[i * ldv + x], acc);
        if (y) acc = b0 * y[x] + acc;
        out[x] = acc * s;
    }
}

// rows j-2 and j-1 of the 1-sync Gram-Schmidt (fgmres.py:60-69), m = j-2 earlier rows:
//   q2 -= sum_i Q[i] s2[i] ; q1 -= sum_i Q[i] s1[i] ; q2 /= norm ; q1 -= q2 r ; q1 /= norm
__global__ void __launch_bounds__(VEC_THREADS) gs1sync_kernel(long long n, int m, const Coefs s2, const Coefs s1, const double* __restrict__ Q,
                                                             long long ldq, double norm, double r, double* q2, double* q1) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long x = blockIdx.x * (long long)blockDim.x + threadIdx.x; x < n; x += stride) {
        double a2 = 0.0, a1 = 0.0;
        for (int i = 0; i < m; ++i) {  // not unrolled: 1.62 ms vs 1.85 ms unrolled by 4 at 20 rows x 53 M values
            const double v = Q[i * ldq + x];
            a2 = fma(v, s2.c[i], a2);
            a1 = fma(v, s1.c[i], a1);
        }
        const double v2 = (q2[x] - a2) / norm;
        q2[x] = v2;
        q1[x] = (q1[x] - a1 - v2 * r) / norm;
    }
}

int grid_for(long long n) {
    long long b = (n + VEC_THREADS - 1) / VEC_THREADS;
    return (int)(b < VEC_BLOCKS ? (b > 0 ? b : 1) : VEC_BLOCKS);
}

}  // namespace

extern "C" {

int wxb_set_expfilter(wxb_handle* h, const double* filt) {
    if (!h) return WXB_EINVAL;
    if (h->cfg.equations != WXB_EQ_EULER3D || !filt) {
        set_err(h, "the exponential filter applies to 3-D Euler handles and needs the (n x n) nodal filter");
        return WXB_EINVAL;
    }
    cudaSetDevice(h->device);
    if (!h->d_expf && cudaMalloc((void**)&h->d_expf, (size_t)h->n * h->n * sizeof(double)) != cudaSuccess) {
        set_err(h, "cudaMalloc of the filter failed");
        return WXB_ENOMEM;
    }
    if (cudaMemcpy(h->d_expf, filt, (size_t)h->n * h->n * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess) {
        set_err(h, "filter upload failed");
        return WXB_ECUDA;
    }
    return WXB_OK;
}

int wxb_expfilter(wxb_handle* h, const double* q, double* out, void* stream) {
    if (!h) return WXB_EINVAL;
    if (!h->d_expf) {
        set_err(h, "wxb_expfilter called before wxb_set_expfilter");
        return WXB_ESTATE;
    }
    if (!q || !out) {
        set_err(h, "null state/output pointer");
        return WXB_EINVAL;
    }
    cudaSetDevice(h->device);
    EvalArgs a{};
    a.q = q;
    a.out = out;
    a.E = h->E;
    a.nz = h->cfg.nz;
    a.nx = h->cfg.nx;
    a.st = h->st;
    cudaStream_t s = (cudaStream_t)stream;
    int r = -1;
    switch (h->n) {
        case 2: r = launch_expfilter<2>(a, h->d_expf, h->d_metric, s); break;
        case 3: r = launch_expfilter<3>(a, h->d_expf, h->d_metric, s); break;
        case 4: r = launch_expfilter<4>(a, h->d_expf, h->d_metric, s); break;
        case 5: r = launch_expfilter<5>(a, h->d_expf, h->d_metric, s); break;
        case 6: r = launch_expfilter<6>(a, h->d_expf, h->d_metric, s); break;
    }
    if (r != 0) {
        set_err(h, "filter kernel launch failed: %s", cudaGetErrorString(cudaGetLastError()));
        return WXB_ECUDA;
    }
    return WXB_OK;
}

int wxb_vec_lincomb3(int64_t n, double a, const double* x, double b, const double* y, double c, const double* z, double dt,
                     double* out, void* stream) {
    if (n < 0 || !x || !z || !out) return WXB_EINVAL;
    if (n == 0) return WXB_OK;
    lincomb3_kernel<<<grid_for(n), VEC_THREADS, 0, (cudaStream_t)stream>>>(n, a, x, b, y, c, z, dt, out);
    return cudaGetLastError() == cudaSuccess ? WXB_OK : WXB_ECUDA;
}

int wxb_vec_nonfinite(int64_t n, const double* x, uint64_t* count_dev, void* stream) {
    if (n < 0 || !x || !count_dev) return WXB_EINVAL;
    cudaStream_t s = (cudaStream_t)stream;
    if (cudaMemsetAsync(count_dev, 0, sizeof(uint64_t), s) != cudaSuccess) return WXB_ECUDA;
    if (n == 0) return WXB_OK;
    nonfinite_kernel<<<grid_for(n), VEC_THREADS, 0, s>>>(n, x, (unsigned long long*)count_dev);
    return cudaGetLastError() == cudaSuccess ? WXB_OK : WXB_ECUDA;
}

int64_t wxb_vec_multidot_scratch(int m, int k) { return (int64_t)VEC_BLOCKS * m * k; }

int wxb_vec_multidot(int64_t n, int m, const double* V, int64_t ldv, int k, const double* W, int64_t ldw, double* out_dev,
                     double* scratch_dev, void* stream) {
    if (n <= 0 || m <= 0 || m > WXB_VEC_MAX_ROWS || (k != 1 && k != 2) || !V || !W || !out_dev || !scratch_dev) return WXB_EINVAL;
    cudaStream_t s = (cudaStream_t)stream;
    // 128-bit loads when every row starts on a 16-byte boundary (WXB_VEC_NOVEC2=1 forces the scalar kernel)
    static const bool novec = getenv("WXB_VEC_NOVEC2") && atoi(getenv("WXB_VEC_NOVEC2"));
    const bool vec2 = !novec && ((uintptr_t)V % 16 == 0) && ((uintptr_t)W % 16 == 0) && (ldv % 2 == 0) && (ldw % 2 == 0 || k == 1);
    const int blocks = grid_for(vec2 ? (n + 1) / 2 : n);
    // more than 8 rows and 128-bit loads possible: all rows (up to 32 per launch) in one pass over the right-hand vectors
    static const bool onepass = !(getenv("WXB_VEC_ONEPASS") && atoi(getenv("WXB_VEC_ONEPASS")) == 0);
    if (onepass && vec2 && m > MD_ROWS) {
        for (int row0 = 0; row0 < m; row0 += 32) {
            const int mm = m - row0 < 32 ? m - row0 : 32;
            const double* Vp = V + row0 * ldv;
#define WXB_MD_ALL(KK, MR) multidot_allrows<KK, MR><<<blocks, VEC_THREADS, 0, s>>>(n, mm, Vp, ldv, W, ldw, scratch_dev, m, row0)
            if (mm <= 16) { if (k == 1) WXB_MD_ALL(1, 16); else WXB_MD_ALL(2, 16); }
            else if (mm <= 24) { if (k == 1) WXB_MD_ALL(1, 24); else WXB_MD_ALL(2, 24); }
            else { if (k == 1) WXB_MD_ALL(1, 32); else WXB_MD_ALL(2, 32); }
#undef WXB_MD_ALL
        }
        multidot_final<<<(m * k + 127) / 128, 128, 0, s>>>(blocks, m * k, scratch_dev, out_dev);
        return cudaGetLastError() == cudaSuccess ? WXB_OK : WXB_ECUDA;
    }
    const int rows = m <= 2 ? 2 : MD_ROWS;
    for (int row0 = 0; row0 < m; row0 += rows) {
        const int mm = m - row0 < rows ? m - row0 : rows;
        const double* Vp = V + row0 * ldv;
#define WXB_MD_LAUNCH(KK, VV, RR) multidot_partial<KK, VV, RR><<<blocks, VEC_THREADS, 0, s>>>(n, mm, Vp, ldv, W, ldw, scratch_dev, m, row0)
        if (rows == 2) {
            if (vec2) { if (k == 1) WXB_MD_LAUNCH(1, 2, 2); else WXB_MD_LAUNCH(2, 2, 2); }
            else      { if (k == 1) WXB_MD_LAUNCH(1, 1, 2); else WXB_MD_LAUNCH(2, 1, 2); }
        } else {
            if (vec2) { if (k == 1) WXB_MD_LAUNCH(1, 2, MD_ROWS); else WXB_MD_LAUNCH(2, 2, MD_ROWS); }
            else      { if (k == 1) WXB_MD_LAUNCH(1, 1, MD_ROWS); else WXB_MD_LAUNCH(2, 1, MD_ROWS); }
        }
#undef WXB_MD_LAUNCH
    }
    multidot_final<<<(m * k + 127) / 128, 128, 0, s>>>(blocks, m * k, scratch_dev, out_dev);
    return cudaGetLastError() == cudaSuccess ? WXB_OK : WXB_ECUDA;
}

// ---- wxb_vec_multidot without any allocation -------------------------------------------------------------------
// Scratch of the two-stage reduction in a file-scope device array (every GPU that loads the library holds these
// WXB_VEC_STATIC_SCRATCH_DOUBLES * 8 bytes = 592 KiB until exit).  In host code `g_md_scratch` is only a stand-in for
// the symbol: the kernels get its device address, resolved once per device with cudaGetSymbolAddress.
#define WXB_VEC_STATIC_SCRATCH_DOUBLES (VEC_BLOCKS * WXB_VEC_STATIC_MAX_ROWS * 2)
static __device__ __align__(256) double g_md_scratch[WXB_VEC_STATIC_SCRATCH_DOUBLES];
static double* g_md_scratch_addr[64];  // per device ordinal, filled at the first call on that device

int wxb_vec_multidot_static(int64_t n, int m, const double* V, int64_t ldv, int k, const double* W, int64_t ldw, double* out_dev,
                            void* stream) {
    if (m > WXB_VEC_STATIC_MAX_ROWS) return WXB_EINVAL;  // beyond the static scratch: use wxb_vec_multidot with caller scratch
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return WXB_ECUDA;
    if (!g_md_scratch_addr[dev]) {  // set-up, once per device, with that device current
        double* ptr = nullptr;
        if (cudaGetSymbolAddress((void**)&ptr, g_md_scratch) != cudaSuccess || !ptr) {
            cudaGetLastError();
            return WXB_ECUDA;
        }
        g_md_scratch_addr[dev] = ptr;
    }
    // every partial sum the final stage reads is written by this call's first stage: nothing has to start at zero
    return wxb_vec_multidot(n, m, V, ldv, k, W, ldw, out_dev, g_md_scratch_addr[dev], stream);
}

int wxb_vec_reduce(int64_t n, int kind, const double* x, const double* y, double* out_dev, double* scratch_dev, void* stream) {
    if (n <= 0 || (kind != 0 && kind != 1) || !x || (kind == 1 && !y) || !out_dev || !scratch_dev) return WXB_EINVAL;
    cudaStream_t s = (cudaStream_t)stream;
    const int blocks = grid_for(n);
    if (kind == 0) {
        reduce_partial<0><<<blocks, VEC_THREADS, 0, s>>>(n, x, y, scratch_dev);
        reduce_final<0><<<1, 32, 0, s>>>(blocks, scratch_dev, out_dev);
    } else {
        reduce_partial<1><<<blocks, VEC_THREADS, 0, s>>>(n, x, y, scratch_dev);
        reduce_final<1><<<1, 32, 0, s>>>(blocks, scratch_dev, out_dev);
    }
    return cudaGetLastError() == cudaSuccess ? WXB_OK : WXB_ECUDA;
}

int wxb_vec_combine(int64_t n, int m, const double* coef_host, const double* V, int64_t ldv, double b0, const double* y, double s,
                    double* out, void* stream) {
    if (n < 0 || m < 0 || m > WXB_VEC_MAX_ROWS || (m > 0 && (!coef_host || !V)) || !out) return WXB_EINVAL;
    if (n == 0) return WXB_OK;
    Coefs cf;
    for (int i = 0; i < m; ++i) cf.c[i] = coef_host[i];
    combine_kernel<<<grid_for(n), VEC_THREADS, 0, (cudaStream_t)stream>>>(n, m, cf, V, ldv, b0, y, s, out);
    return cudaGetLastError() == cudaSuccess ? WXB_OK : WXB_ECUDA;
}

int wxb_vec_gs1sync(int64_t n, int m, const double* s2_host, const double* s1_host, const double* Q, int64_t ldq, double norm,
                    double r, double* q2, double* q1, void* stream) {
    if (n <= 0 || m < 0 || m > WXB_VEC_MAX_ROWS || (m > 0 && (!s2_host || !s1_host || !Q)) || !q2 || !q1 || norm == 0.0) return WXB_EINVAL;
    Coefs c2, c1;
    for (int i = 0; i < m; ++i) {
        c2.c[i] = s2_host[i];
        c1.c[i] = s1_host[i];
    }
    gs1sync_kernel<<<grid_for(n), VEC_THREADS, 0, (cudaStream_t)stream>>>(n, m, c2, c1, Q, ldq, norm, r, q2, q1);
    return cudaGetLastError() == cudaSuccess ? WXB_OK : WXB_ECUDA;
}

}  // extern "C"
