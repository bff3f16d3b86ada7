// 3-D Euler DFR right-hand side on the cubed sphere -- sm_100a kernels (FP64, CUDA-core FMA).
//
// Two kernels per evaluation instead of the reference's ~150 numpy expressions:
//
//   euler_traces  element -> 6 face traces (rho and rho*theta extrapolated in log space), stored in the form the flux
//                 consumes: rho, u1, u2, w, rho*theta, sound speed (WXB_TRACE_FIELDS)
//                 = RHSDirecFluxReconstruction_mpi.solution_extrapolation (wx_factory/rhs/rhs_dfr.py:50-71)
//   euler_main    everything else, one thread per solution point, element fields staged in shared memory:
//                 pointwise fluxes          wx_factory/pde/pde_euler_cubesphere.py:72-124
//                 interior derivative       wx_factory/rhs/rhs_dfr.py:25-34,89-104 (sum-factorised: 2n flops
//                                           per value instead of the dense kron GEMM's 2n^3)
//                 halo "receive"            wx_factory/rhs/rhs_dfr.py:174-201 as a table-driven pull of the
//                                           neighbour's trace incl. the panel-edge rotation/flip of
//                                           wx_factory/process_topology.py:128-175
//                 vertical BC + Rusanov     wx_factory/rhs/rhs_dfr.py:265-268,
//                                           wx_factory/pde/pde_euler_cubesphere.py:145-160, wx_factory/pde/fluxes.py:150-222,326-403,507-582
//                 lift + split rho-w form   wx_factory/rhs/rhs_dfr.py:106-139
//                 forcing                   wx_factory/pde/pde_euler_cubesphere.py:203-290, wx_factory/init/dcmip.py:676-755
//   the FD Jacobian-vector product (wx_factory/solvers/matvec.py:57-88) is the same pair with the perturbed
//   state q + eps*v formed in registers and dt*(. - rhs0)/eps applied in the epilogue.
//
// Local point orders (wx_factory/geometry/cubed_sphere_3d.py:764-775):
//   node = k*n^2 + j*n + i ; x1-face pt = k*n + j ; x2-face pt = k*n + i ; x3-face pt = j*n + i
//   faces: 0 W(x1-) 1 E(x1+) 2 S(x2-) 3 N(x2+) 4 D(x3-) 5 U(x3+)
#include "fastmath.cuh"
#include "wxb_internal.cuh"

// tuning knobs per order n: elements per block (EB) and resident blocks per SM the main kernel is compiled for (MINB);
// the defaults are the best of the sweeps recorded in profiles/ (scripts/gpu_variants.sh)
#ifndef WXB_EB2
#define WXB_EB2 8
#endif
#ifndef WXB_EB3
#define WXB_EB3 4
#endif
#ifndef WXB_EB4
#define WXB_EB4 1
#endif
#ifndef WXB_EB5
#define WXB_EB5 1
#endif
#ifndef WXB_EB6
#define WXB_EB6 1
#endif
#ifndef WXB_MINB2
#define WXB_MINB2 8
#endif
#ifndef WXB_MINB3
#define WXB_MINB3 4
#endif
#ifndef WXB_MINB4
#define WXB_MINB4 7  // with the (rho, u, rho theta, a) traces: 7-8 blocks (122-142 registers, no spills) beat 10 (96, spills)
#endif
#ifndef WXB_MINB5
#define WXB_MINB5 4
#endif
#ifndef WXB_MINB6
#define WXB_MINB6 3
#endif
#ifndef WXB_CH3
#define WXB_CH3 0  // 1: Christoffel symbols loaded 9 at a time (fewer registers, less memory parallelism)
#endif
#ifndef WXB_FR_MIN_N
#define WXB_FR_MIN_N 4  // lowest order that builds the per-face address table in euler_main
#endif
#ifndef WXB_RCP
#define WXB_RCP 1  // velocities through one reciprocal of rho per point instead of three divisions
#endif
#ifndef WXB_FASTMATH
#define WXB_FASTMATH 0  // 1: table-driven exp / log / reciprocal / sqrt of fastmath.cuh; here the tables are read through L1 from
                        // global memory and that costs 2-3 % (A/B on B200, profiles/r02_libs_ab.log): off on this path
#endif

// Formulation knobs of the trace / interface path (A/B'd on B200: profiles/r01_trace_variants.log):
#ifndef WXB_TRACE_FIELDS
#define WXB_TRACE_FIELDS 1  // 0: rho, rho u1, rho u2, rho w, rho theta, log(rho theta)  (face pressure = one exp per side and reader)
                            // 1: rho, u1, u2, w, rho theta, a = sqrt(gamma p / rho)      (no exp / reciprocal in the interface pass)
                            // 2: rho, rho u1, rho u2, rho w, rho theta, p                (no exp in the interface pass)
#endif
#ifndef WXB_FOLD_LOGP
#define WXB_FOLD_LOGP 1     // derivative + lift of log p as one matrix on log(Rd/p0 rho theta) (see O_DLP); needed by fields 1, 2
#endif
#ifndef WXB_TRACE_MAP
#define WXB_TRACE_MAP 0     // euler_traces: 0 = both sides of a line per thread (shares the line's loads), 1 = one (line, side)
                            // per thread and round (all lanes busy, but 0.32 vs 0.22 ms: two dependent rounds on warp 0)
#endif
#ifndef WXB_SYM
#define WXB_SYM 1           // fields 1: Rusanov flux written in (own, facing) form without sorting the states into L / R
#endif
// elements per block of euler_traces; its lines are spread over ALL threads of the block.  With one 64-thread element per
// block at n = 4 the 48 lines fill one warp and half of the other; 4 elements per block fill 6 warps of 8 completely but
// measured slower on B200 (traces 0.251 vs 0.230 ms, profiles/r01_variants_teb.log): kept at the main kernel's values
#ifndef WXB_TEB2
#define WXB_TEB2 8
#endif
#ifndef WXB_TEB3
#define WXB_TEB3 4
#endif
#ifndef WXB_TEB4
#define WXB_TEB4 1
#endif
#ifndef WXB_TEB5
#define WXB_TEB5 1
#endif
#ifndef WXB_TEB6
#define WXB_TEB6 1
#endif
#ifndef WXB_TR_MINB
#define WXB_TR_MINB 0       // resident blocks per SM euler_traces is compiled for (0: no register cap)
#endif
#if WXB_TRACE_FIELDS != 0 && !WXB_FOLD_LOGP
#error "trace fields 1 and 2 carry no log(rho theta): they need WXB_FOLD_LOGP"
#endif

namespace {

// exp / log / 1/x / sqrt: fastmath.cuh versions (tab = EvalArgs::fasttab in global memory, 4.6 KB that stay in L1) or
// the CUDA library ones
__device__ __forceinline__ double m_exp(double x, const double* __restrict__ tab) {
    return WXB_FASTMATH ? wxbfm::fast_exp(x, tab + WXB_EXPTAB_OFF) : exp(x);
}
__device__ __forceinline__ double m_log(double x, const double* __restrict__ tab) { return WXB_FASTMATH ? wxbfm::fast_log(x, tab) : log(x); }
__device__ __forceinline__ double m_rcp(double x) { return WXB_FASTMATH ? wxbfm::fast_rcp(x) : 1.0 / x; }
__device__ __forceinline__ double m_sqrt(double x) { return WXB_FASTMATH ? wxbfm::fast_sqrt(x) : sqrt(x); }

constexpr int NVAR = 5;
constexpr int NTR = 6;                      // trace fields per face point (WXB_TRACE_FIELDS)
constexpr int NFF = WXB_FOLD_LOGP ? 6 : 7;  // interface results per face point kept for the lift

// extrapolated line sums -> the trace fields (l0 = log rho, l4 = log(rho theta), m1..m3 = momenta)
__device__ __forceinline__ void trace_fields(double m1, double m2, double m3, double l0, double l4, const double* __restrict__ tab, double out[NTR]) {
    const double rho = m_exp(l0, tab);
    out[0] = rho;
    out[4] = m_exp(l4, tab);
#if WXB_TRACE_FIELDS == 1
    const double r = m_rcp(rho);
    out[1] = m1 * r;
    out[2] = m2 * r;
    out[3] = m3 * r;
    // a = sqrt(gamma p / rho) with p = p0*exp(gamma*log(Rd/p0 * rho_theta)) (pde_euler_cubesphere.py:158-160,
    // fluxes.py:171-172) from the extrapolated logs with one exp and no sqrt: sqrt(gamma p0) exp((gamma(l4+c) - l0)/2)
    out[5] = WXB_SQRT_GAMMA_P0 * m_exp(0.5 * (WXB_GAMMA * (l4 + WXB_LOG_RD_P0) - l0), tab);
#else
    out[1] = m1;
    out[2] = m2;
    out[3] = m3;
#if WXB_TRACE_FIELDS == 2
    out[5] = WXB_P0 * m_exp(WXB_GAMMA * (l4 + WXB_LOG_RD_P0), tab);  // pde_euler_cubesphere.py:158-160 from the extrapolated log
#else
    out[5] = l4;
#endif
#endif
}

template <int N>
struct Cfg {
    static constexpr int P = N * N * N;
    static constexpr int F = N * N;
    // elements per block: aim for ~256 threads
    static constexpr int EB = (N == 2) ? WXB_EB2 : (N == 3) ? WXB_EB3 : (N == 4) ? WXB_EB4 : (N == 5) ? WXB_EB5 : WXB_EB6;
    static constexpr int THREADS = EB * P;
    static constexpr int TEB = (N == 2) ? WXB_TEB2 : (N == 3) ? WXB_TEB3 : (N == 4) ? WXB_TEB4 : (N == 5) ? WXB_TEB5 : WXB_TEB6;  // euler_traces
    static constexpr int TTHREADS = TEB * P;
    // resident blocks per SM the main kernel is compiled for (register budget = 65536 / (THREADS * MINB))
    static constexpr int MINB = (N == 2) ? WXB_MINB2 : (N == 3) ? WXB_MINB3 : (N == 4) ? WXB_MINB4 : (N == 5) ? WXB_MINB5 : WXB_MINB6;
    // offsets into the operator block
    static constexpr int O_DIFF = 2 * N, O_CORR = 2 * N + N * N, O_HF = 4 * N + N * N;
    // derivative + lift of log p as ONE matrix acting on lg = log(Rd/p0 * rho theta) of the line, and the constant
    // left over (built in extended precision at create, wxb_api.cu)
    static constexpr int O_DLP = 4 * N + 2 * N * N, O_KLP = O_DLP + N * N;
    static constexpr int NOPS = 5 * N + 3 * N * N;
};

__device__ __forceinline__ double pressure_of(double rho_theta) {
    // p0 * exp((cpd/cvd) * log((Rd/p0) * rho_theta))   (pde_euler_cubesphere.py:100)
    return WXB_P0 * exp(WXB_GAMMA * log((WXB_RD / WXB_P0) * rho_theta));
}

__device__ __forceinline__ int element_of(const EvalArgs& a, int slot) {
    if (a.elist) return slot < a.nelist ? a.elist[slot] : -1;
    return slot < a.E ? slot : -1;
}

// ------------------------------------------------------------------------------------------------------
// Element -> face extrapolation of one line of N solution points (rhs_dfr.py:50-71, operators.py:161-165).
// rows: the element's q rows 1..3 (momenta) and the rows of log(rho), log(rho theta), each P doubles, `rs` apart.
// Returns the 6 trace fields: rho, rho u1, rho u2, rho w, rho theta, log(rho theta); rho and rho theta are
// extrapolated in log space.  euler_traces (what the neighbours read) and euler_main (the element's own side)
// both call this, so the two sides of an interface see bit-identical values.
template <int N>
__device__ __forceinline__ void extrap_line(const double* __restrict__ ext, const double* mom, int mom_rs, const double* lrho,
                                            const double* lrt, int base, int step, const double* __restrict__ tab, double out[NTR]) {
    double a1 = 0.0, a2 = 0.0, a3 = 0.0, l0 = 0.0, l4 = 0.0;
#pragma unroll
    for (int m = 0; m < N; ++m) {
        const double w = ext[m];
        const int o = base + m * step;
        a1 = fma(w, mom[o], a1);
        a2 = fma(w, mom[mom_rs + o], a2);
        a3 = fma(w, mom[2 * mom_rs + o], a3);
        l0 = fma(w, lrho[o], l0);
        l4 = fma(w, lrt[o], l4);
    }
    trace_fields(a1, a2, a3, l0, l4, tab, out);
}

// both sides of one line at once (shares the loads); per side the same fma chain as extrap_line
template <int N>
__device__ __forceinline__ void extrap_line2(const double* __restrict__ ext, const double* mom, int mom_rs, const double* lrho,
                                             const double* lrt, int base, int step, const double* __restrict__ tab, double lo[NTR], double hi[NTR]) {
    double a[5] = {0, 0, 0, 0, 0}, b[5] = {0, 0, 0, 0, 0};
#pragma unroll
    for (int m = 0; m < N; ++m) {
        const double wl = ext[m], wh = ext[N + m];
        const int o = base + m * step;
        const double x[5] = {mom[o], mom[mom_rs + o], mom[2 * mom_rs + o], lrho[o], lrt[o]};
#pragma unroll
        for (int v = 0; v < 5; ++v) {
            a[v] = fma(wl, x[v], a[v]);
            b[v] = fma(wh, x[v], b[v]);
        }
    }
    trace_fields(a[0], a[1], a[2], a[3], a[4], tab, lo);
    trace_fields(b[0], b[1], b[2], b[3], b[4], tab, hi);
}

// line of face point pt of direction d: first node and node stride
template <int N>
__device__ __forceinline__ void line_of(int d, int pt, int& base, int& step) {
    const int a1 = pt / N, a0 = pt % N;
    if (d == 0) { base = a1 * N * N + a0 * N; step = 1; }       // pt=(k,j): line along i
    else if (d == 1) { base = a1 * N * N + a0; step = N; }      // pt=(k,i): line along j
    else { base = a1 * N + a0; step = N * N; }                  // pt=(j,i): line along k
}

// ------------------------------------------------------------------------------------------------------
// traces: every element's 6 faces x 6 fields -> a.trace (read by the neighbours in euler_main, packed for other ranks)
template <int N>
__global__ void
#if WXB_TR_MINB
__launch_bounds__(Cfg<N>::TTHREADS, WXB_TR_MINB)
#else
__launch_bounds__(Cfg<N>::TTHREADS)
#endif
euler_traces(const EvalArgs a) {
    using C = Cfg<N>;
    constexpr int P = C::P, F = C::F, EB = C::TEB;
    __shared__ double s_ext[2 * N];
    __shared__ double s_q[EB][5][P];  // rho u1, rho u2, rho w, log(rho), log(rho theta)
    __shared__ int s_e[EB];

    const int le = threadIdx.x / P, node = threadIdx.x % P;
    const int e = element_of(a, blockIdx.x * EB + le);  // all elements, or those of a.elist (chunked host path)
    const bool active = e >= 0;
    if (threadIdx.x < 2 * N) s_ext[threadIdx.x] = a.ops[threadIdx.x];
    if (node == 0) s_e[le] = e;
    if (active) {
        const ElemPos ep = decode_elem(a.st, e, a.nz, a.nx, P, NVAR);
        double x[NVAR];  // all loads first: the conditional store below must not fence them
#pragma unroll
        for (int v = 0; v < NVAR; ++v) x[v] = a.q[ep.qoff + v * ep.vstride + node];
        if (a.v) {  // the perturbed state of the FD Jacobian-vector product, kept for euler_main
            double dv[NVAR];
#pragma unroll
            for (int v = 0; v < NVAR; ++v) dv[v] = a.v[ep.qoff + v * ep.vstride + node];
#pragma unroll
            for (int v = 0; v < NVAR; ++v) {
                x[v] += a.eps * dv[v];
                a.qp[ep.qoff + v * ep.vstride + node] = x[v];
            }
        }
        s_q[le][0][node] = x[1];
        s_q[le][1][node] = x[2];
        s_q[le][2][node] = x[3];
        s_q[le][3][node] = m_log(x[0], a.fasttab);
        s_q[le][4][node] = m_log(x[4], a.fasttab);
    }
    __syncthreads();
#if WXB_TRACE_MAP
    if (!active) return;
    // 6F trace points per element, one (line, side) per thread and round: all lanes of a warp carry a point (with both
    // sides per thread the 3F lines fill only 3/4 of the block's lanes, and the per-point exp/reciprocal work dominates)
    for (int tp = node; tp < 6 * F; tp += P) {
        const int f = tp / F, pt = tp % F, d = f >> 1, side = f & 1;
        int base, step;
        line_of<N>(d, pt, base, step);
        double o[NTR];
        extrap_line<N>(s_ext + side * N, &s_q[le][0][0], P, &s_q[le][3][0], &s_q[le][4][0], base, step, a.fasttab, o);
        double* t = a.trace + (((long long)e * 6 + f) * NTR) * F + pt;  // [E][face][field][F]
#pragma unroll
        for (int v = 0; v < NTR; ++v) t[v * F] = o[v];
    }
#else
    // 3F lines per element, each yields the min-side and the max-side face value (same arithmetic per side as
    // extrap_line, which the TMA-staged euler_main uses for the element's own side); the block's EB*3F lines are spread
    // over all its threads
    for (int lt = threadIdx.x; lt < EB * 3 * F; lt += EB * P) {
        const int el = lt / (3 * F), ln = lt % (3 * F);
        const int e2 = s_e[el];
        if (e2 < 0) continue;
        const int d = ln / F, pt = ln % F;
        int base, step;
        line_of<N>(d, pt, base, step);
        double lo[NTR], hi[NTR];
        extrap_line2<N>(s_ext, &s_q[el][0][0], P, &s_q[el][3][0], &s_q[el][4][0], base, step, a.fasttab, lo, hi);
        double* t = a.trace + (((long long)e2 * 6 + 2 * d) * NTR) * F + pt;  // [E][face][field][F]
#pragma unroll
        for (int v = 0; v < NTR; ++v) {
            t[v * F] = lo[v];
            t[(NTR + v) * F] = hi[v];
        }
    }
#endif
}


// ------------------------------------------------------------------------------------------------------
// Per-element bookkeeping, computed once per element by one thread and kept in shared memory so that the
// hot loops carry no index decoding.
struct ElemInfo {
    long long qoff, vstride;   // (var 0, this element, node 0) in the state; distance between variables
    const double* fm[3];       // (comp 0, min-side face, pt 0) of this element in fm_i / fm_j / fm_k
    long long fm_side[3];      // min-side -> max-side face
    int e, k, c, ny;
};

__device__ __forceinline__ void make_info(const EvalArgs& a, int e, int F, int P, ElemInfo& o) {
    const ElemPos ep = decode_elem(a.st, e, a.nz, a.nx, P, NVAR);
    o.qoff = ep.qoff;
    o.vstride = ep.vstride;
    o.e = e;
    o.k = ep.k;
    o.c = ep.c;
    o.ny = ep.ny;
    const long long nz = a.nz, ny = ep.ny, nx = a.nx;
    // interface metric, face-major per slab: [face][4 components][F] (transposed at create, wxb_api.cu)
    o.fm[0] = a.fmi + a.st.fmi[ep.s] + (((long long)ep.k * ny + ep.jl) * (nx + 1) + ep.i) * (4 * F);
    o.fm_side[0] = 4 * F;
    o.fm[1] = a.fmj + a.st.fmj[ep.s] + (((long long)ep.k * (ny + 1) + ep.jl) * nx + ep.i) * (4 * F);
    o.fm_side[1] = nx * (4 * F);
    o.fm[2] = a.fmk + a.st.fmk[ep.s] + (((long long)ep.k * ny + ep.jl) * nx + ep.i) * (4 * F);
    o.fm_side[2] = ny * nx * (4 * F);
    (void)nz;
}

// ------------------------------------------------------------------------------------------------------
// one face point: common Rusanov flux and the own-side split rho-w pieces
//   own / nb : the 6 trace fields of this element's side and of the facing side (already flipped / rotated)
//   out[0..3] = F* rows rho, rho u1, rho u2, rho theta ; out[4] = wflux_adv* ; out[5] = wflux_pres* (own side) ;
//   out[6] = log(p_itf own side)                 (fluxes.py:150-222,326-403,507-582; rhs_dfr.py:106-139)
#if WXB_TRACE_FIELDS == 1
// traces carry (rho, u1, u2, w, rho theta, a): momenta and pressure are products, the signal speed needs one sqrt of
// the metric term; out[0..5] only (the own-side log p of the lift is folded into the derivative matrix, see O_DLP)
template <int N>
__device__ __forceinline__ void rusanov_point(const EvalArgs& a, const ElemInfo& ei, int d, int side, int pt, bool bc,
                                              const double own[NTR], const double nb[NTR], const double fm[4], double out[NFF]) {
    const double sg = fm[0], hd0 = fm[1], hd1 = fm[2], hd2 = fm[3];
    const double hdd = d == 0 ? hd0 : (d == 1 ? hd1 : hd2);
#if WXB_SYM
    // The Rusanov flux is symmetric in its two states up to the sign of the jump: with o = own side, f = facing side
    // and sgn = +1 when the own side is L (max-side face), F* = (F(o) + F(f))/2 - sgn * eig sqrtG (q_f - q_o)/2.
    // No L/R sorting of the 12 values.
    const double uo = d == 0 ? own[1] : (d == 1 ? own[2] : own[3]);
    double uf = d == 0 ? nb[1] : (d == 1 ? nb[2] : nb[3]);
    if (bc) uf = -uo;  // odd symmetry of w across the rigid boundary (pde_euler_cubesphere.py:152-156)
    constexpr double inv_gamma = 1.0 / WXB_GAMMA;
    const double po = own[5] * own[5] * (own[0] * inv_gamma), pf = nb[5] * nb[5] * (nb[0] * inv_gamma);  // a^2 rho / gamma
    const double sh = m_sqrt(hdd);
    const double eig = fmax(fabs(uo) + sh * own[5], fabs(uf) + sh * nb[5]);  // |u| + sqrt(h^dd gamma p / rho)
    const double sguo = sg * uo, sguf = sg * uf, es = side ? eig * sg : -(eig * sg);
    const double mo1 = own[0] * own[1], mo2 = own[0] * own[2], mo3 = own[0] * own[3];
    const double mf1 = nb[0] * nb[1], mf2 = nb[0] * nb[2], mf3 = nb[0] * nb[3];
    out[0] = 0.5 * (sguo * own[0] + sguf * nb[0] - es * (nb[0] - own[0]));
    out[1] = 0.5 * ((sguo * mo1 + sg * hd0 * po) + (sguf * mf1 + sg * hd0 * pf) - es * (mf1 - mo1));
    out[2] = 0.5 * ((sguo * mo2 + sg * hd1 * po) + (sguf * mf2 + sg * hd1 * pf) - es * (mf2 - mo2));
    out[3] = 0.5 * (sguo * own[4] + sguf * nb[4] - es * (nb[4] - own[4]));
    out[4] = 0.5 * (sguo * mo3 + sguf * mf3 - es * (mf3 - mo3));
    out[5] = 0.5 * (sg * hd2 * po + sg * hd2 * pf) * m_rcp(po);
}
#else
    double L[NTR], R[NTR];
#pragma unroll
    for (int v = 0; v < NTR; ++v) {
        L[v] = side ? own[v] : nb[v];
        R[v] = side ? nb[v] : own[v];
    }
    double uL = d == 0 ? L[1] : (d == 1 ? L[2] : L[3]);
    double uR = d == 0 ? R[1] : (d == 1 ? R[2] : R[3]);
    if (bc) {  // odd symmetry of w across the rigid boundary (pde_euler_cubesphere.py:152-156)
        if (side == 0) uL = -uR; else uR = -uL;
    }
    constexpr double inv_gamma = 1.0 / WXB_GAMMA;
    const double pL = L[5] * L[5] * (L[0] * inv_gamma), pR = R[5] * R[5] * (R[0] * inv_gamma);  // a^2 rho / gamma
    const double sh = sqrt(hdd);
    const double eig = fmax(fabs(uL) + sh * L[5], fabs(uR) + sh * R[5]);  // |u| + sqrt(h^dd gamma p / rho)
    const double sguL = sg * uL, sguR = sg * uR, es = eig * sg;
    const double mL1 = L[0] * L[1], mL2 = L[0] * L[2], mL3 = L[0] * L[3];
    const double mR1 = R[0] * R[1], mR2 = R[0] * R[2], mR3 = R[0] * R[3];
    out[0] = 0.5 * (sguL * L[0] + sguR * R[0] - es * (R[0] - L[0]));
    out[1] = 0.5 * ((sguL * mL1 + sg * hd0 * pL) + (sguR * mR1 + sg * hd0 * pR) - es * (mR1 - mL1));
    out[2] = 0.5 * ((sguL * mL2 + sg * hd1 * pL) + (sguR * mR2 + sg * hd1 * pR) - es * (mR2 - mL2));
    out[3] = 0.5 * (sguL * L[4] + sguR * R[4] - es * (R[4] - L[4]));
    out[4] = 0.5 * (sguL * mL3 + sguR * mR3 - es * (mR3 - mL3));
    const double pown = side ? pL : pR;
    out[5] = 0.5 * (sg * hd2 * pL + sg * hd2 * pR) / pown;
}
#endif
#else
template <int N>
__device__ __forceinline__ void rusanov_point(const EvalArgs& a, const ElemInfo& ei, int d, int side, int pt, bool bc,
                                              const double own[NTR], const double nb[NTR], const double fm[4], double out7[NFF]) {
    // interface metric, single-valued per physical face
    const double sg = fm[0], hd0 = fm[1], hd1 = fm[2], hd2 = fm[3];
    const double hdd = d == 0 ? hd0 : (d == 1 ? hd1 : hd2);
    // L = max-side face of the lower element, R = min-side face of the upper element
    double L[NTR], R[NTR];
#pragma unroll
    for (int v = 0; v < NTR; ++v) {
        L[v] = side ? own[v] : nb[v];
        R[v] = side ? nb[v] : own[v];
    }
    const double vL = d == 0 ? L[1] : (d == 1 ? L[2] : L[3]);
    const double vR = d == 0 ? R[1] : (d == 1 ? R[2] : R[3]);
#if WXB_RCP
    const double rL = 1.0 / L[0], rR = 1.0 / R[0];  // one reciprocal per side serves the velocity and the sound speed
    double uL = vL * rL;
    double uR = vR * rR;
#else
    double uL = vL / L[0];
    double uR = vR / R[0];
#endif
    if (bc) {  // odd symmetry of w across the rigid boundary (pde_euler_cubesphere.py:152-156)
        if (side == 0) uL = -uR; else uR = -uL;
    }
#if WXB_TRACE_FIELDS == 2
    const double pL = L[5], pR = R[5];  // the face pressure comes with the trace
#else
    // face pressure p0*exp(gamma*log(Rd/p0 * rho_theta)) (pde_euler_cubesphere.py:158-160) with
    // log(rho_theta) taken from the trace (field 5) instead of log(exp(.)): same value to rounding
    const double pL = WXB_P0 * exp(WXB_GAMMA * (L[5] + WXB_LOG_RD_P0));
    const double pR = WXB_P0 * exp(WXB_GAMMA * (R[5] + WXB_LOG_RD_P0));
#endif
#if WXB_RCP
    const double eigL = fabs(uL) + sqrt(hdd * WXB_GAMMA * pL * rL);
    const double eigR = fabs(uR) + sqrt(hdd * WXB_GAMMA * pR * rR);
#else
    const double eigL = fabs(uL) + sqrt(hdd * WXB_GAMMA * pL / L[0]);
    const double eigR = fabs(uR) + sqrt(hdd * WXB_GAMMA * pR / R[0]);
#endif
    const double eig = fmax(eigL, eigR);
    const double sguL = sg * uL, sguR = sg * uR, es = eig * sg;
    out7[0] = 0.5 * (sguL * L[0] + sguR * R[0] - es * (R[0] - L[0]));
    out7[1] = 0.5 * ((sguL * L[1] + sg * hd0 * pL) + (sguR * R[1] + sg * hd0 * pR) - es * (R[1] - L[1]));
    out7[2] = 0.5 * ((sguL * L[2] + sg * hd1 * pL) + (sguR * R[2] + sg * hd1 * pR) - es * (R[2] - L[2]));
    out7[3] = 0.5 * (sguL * L[4] + sguR * R[4] - es * (R[4] - L[4]));
    out7[4] = 0.5 * (sguL * L[3] + sguR * R[3] - es * (R[3] - L[3]));
    const double pown = side ? pL : pR;
    out7[5] = 0.5 * (sg * hd2 * pL + sg * hd2 * pR) / pown;
#if !WXB_FOLD_LOGP
    const double lpL = WXB_LOG_P0 + WXB_GAMMA * (L[5] + WXB_LOG_RD_P0);
    const double lpR = WXB_LOG_P0 + WXB_GAMMA * (R[5] + WXB_LOG_RD_P0);
    out7[6] = side ? lpL : lpR;  // log(p_itf) of the own side (rhs_dfr.py:113)
#endif
}

#endif

// where the facing side's trace row (field 0, level of this element) starts, in units of F doubles; the distance
// between fields; flags (bit0 = reversed along the edge, bits 8.. = 1 + xform row).  Vertical faces have no table.
struct NbrRow {
    long long row0, vs;
    int flags;
    bool bc;
};
__device__ __forceinline__ NbrRow nbr_row(const EvalArgs& a, int e, int k, int c, int ny, int f) {
    NbrRow r;
    r.flags = 0;
    r.bc = false;
    if (f >= 4) {
        r.bc = (f == 4) ? (k == 0) : (k == a.nz - 1);
        const int e2 = (f == 4) ? e - ny * a.nx : e + ny * a.nx;
        r.row0 = ((long long)e2 * 6 + (f ^ 1)) * NTR;
        r.vs = 1;
    } else {
        const int4 t = a.nbr[(long long)f * a.C + c];
        r.row0 = (long long)t.x + (long long)k * t.y;
        r.vs = t.z;
        r.flags = t.w;
    }
    return r;
}

// Everything the interface pass needs to address one face of one element, worked out once by one thread so that
// the 6F face points only add their point index.
struct FaceRow {
    const double* own;  // trace field 0 of this face, point 0
    const double* nb;   // facing side's field 0, point 0 (the own row again at a rigid boundary: mirrored ghost)
    const double* fm;   // interface metric, component 0, point 0
    long long nb_vs;    // distance between the facing row's fields
    int flags, bc;
};
__device__ __forceinline__ void make_face_row(const EvalArgs& a, int e, const ElemPos& ep, int f, int F, FaceRow& o) {
    const NbrRow r = nbr_row(a, e, ep.k, ep.c, ep.ny, f);
    const int d = f >> 1, side = f & 1;
    const long long nz = a.nz, ny = ep.ny, nx = a.nx;
    o.own = a.trace + (((long long)e * 6 + f) * NTR) * F;
    o.nb = r.bc ? o.own : a.trace + r.row0 * F;
    o.nb_vs = F;  // element-major trace layout: fields of a face are always one row apart (checked at create)
    o.flags = r.flags;
    o.bc = r.bc;
    if (d == 0) {
        o.fm = a.fmi + a.st.fmi[ep.s] + (((long long)ep.k * ny + ep.jl) * (nx + 1) + ep.i + side) * (4 * F);
    } else if (d == 1) {
        o.fm = a.fmj + a.st.fmj[ep.s] + (((long long)ep.k * (ny + 1) + ep.jl + side) * nx + ep.i) * (4 * F);
    } else {
        o.fm = a.fmk + a.st.fmk[ep.s] + ((((long long)ep.k + side) * ny + ep.jl) * nx + ep.i) * (4 * F);
    }
}

// reorder / rotate the facing side's values into this element's frame (process_topology.py:128-175)
template <int N>
__device__ __forceinline__ void apply_xform(const EvalArgs& a, int flags, int al, double nb[NTR]) {
    const int xid = flags >> 8;
    if (xid) {
        const double* m = a.xform + ((long long)(xid - 1) * N + al) * 4;
        const double b1 = m[0] * nb[1] + m[1] * nb[2];
        const double b2 = m[2] * nb[1] + m[3] * nb[2];
        nb[1] = b1;
        nb[2] = b2;
    }
}

// ------------------------------------------------------------------------------------------------------
// Rows of the per-element shared-memory area (each P doubles): 0-17 flux fields (d*6 + {F_rho, F_u1, F_u2, F_theta,
// w_adv, w_pres}), 18 = log(Rd/p0 rho theta), 19 = sqrtG*rho.  A thread only ever writes its own node's column.
constexpr int NFLUX = 20, RLOGP = 18, RSGRHO = 19;
constexpr int R_CH = 0, R_SG = 27, R_H = 28, R_IDZ = 34;  // rows of the element's metric block [35][P] (wxb_api.cu)
constexpr int NMETRIC = 35;

template <int N>
struct MainSmem {
    alignas(16) double st[NFLUX][Cfg<N>::EB][Cfg<N>::P];  // row-major across the block's elements
    double ff[Cfg<N>::EB][6 * NFF * Cfg<N>::F];            // interface results (face*NFF + value, point)
    double op[Cfg<N>::NOPS];
    ElemInfo ei[Cfg<N>::EB];
    NbrRow nr[Cfg<N>::EB][6];
    FaceRow fr[Cfg<N>::EB][6];
};

// ------------------------------------------------------------------------------------------------------
// main kernel: one thread per solution point, EB elements per block.
//   interface pass  per face point: own trace, facing trace (pull table), Rusanov flux -> shared memory
//   node pass       per node: forcing, the flux fields of the three directions -> shared memory
//   assemble        per node: sum-factorised derivative + lift, epilogue
// In JVP mode the state rows come from the perturbed state q + eps*v written by euler_traces (a.qp).
template <int N, bool JVP>
__global__ void __launch_bounds__(Cfg<N>::THREADS, Cfg<N>::MINB) euler_main(const EvalArgs a) {
    using C = Cfg<N>;
    constexpr int P = C::P, F = C::F, EB = C::EB;
    constexpr bool USE_FR = N >= WXB_FR_MIN_N;  // per-face address table
    extern __shared__ __align__(16) unsigned char smem_raw[];
    using SM = MainSmem<N>;
    SM& sm = *reinterpret_cast<SM*>(smem_raw);
    double* const s_op = sm.op;

    const int le = threadIdx.x / P, node = threadIdx.x % P;
    const int e = element_of(a, blockIdx.x * EB + le);
    const bool active = e >= 0;
    const double* const qsrc = JVP ? a.qp : a.q;
    const long long EP = (long long)a.E * P;
    for (int t = threadIdx.x; t < C::NOPS; t += C::THREADS) s_op[t] = a.ops[t];
    if (active && node == 0) make_info(a, e, F, P, sm.ei[le]);
    if (active && node >= P - 6) {  // six other threads look the facing rows up (one table load each)
        const ElemPos ep = decode_elem(a.st, e, a.nz, a.nx, P, NVAR);
        if (!USE_FR) sm.nr[le][node - (P - 6)] = nbr_row(a, e, ep.k, ep.c, ep.ny, node - (P - 6));
        else make_face_row(a, e, ep, node - (P - 6), F, sm.fr[le][node - (P - 6)]);
    }
    __syncthreads();
    constexpr int RS = EB * P;  // distance between rows of one element
    auto ffp = [&](int fv, int pt) -> double* { return &sm.ff[le][fv * F + pt]; };

    // ---- interface fluxes, 6F face points per element (their registers die before the node pass) ----
    if (active) {
        const ElemInfo& ei = sm.ei[le];
        for (int fp = node; fp < 6 * F; fp += P) {
            const int f = fp / F, pt = fp % F, d = f >> 1, side = f & 1;
            const int kl = pt / N, al = pt % N;
            double own[NTR], nb[NTR], fm[4];
            int flags;
            bool bc;
            if (!USE_FR) {  // low orders: a face table per element costs more than it saves (few points per face)
                const NbrRow& nr = sm.nr[le][f];
                bc = nr.bc;
                flags = nr.flags;
                const int pt2 = (flags & 1) ? kl * N + (N - 1 - al) : pt;
                const double* m = ei.fm[d] + (side ? ei.fm_side[d] : 0) + pt;
                fm[0] = m[0]; fm[1] = m[F]; fm[2] = m[2 * F]; fm[3] = m[3 * F];
                const double* t = a.trace + (((long long)ei.e * 6 + f) * NTR) * F + pt;
                const double* src = bc ? t : a.trace + nr.row0 * F + pt2;  // mirrored ghost (rhs_dfr.py:265-268)
#pragma unroll
                for (int v = 0; v < NTR; ++v) own[v] = t[v * F];
#pragma unroll
                for (int v = 0; v < NTR; ++v) nb[v] = src[v * F];
            } else {
                const FaceRow& fr = sm.fr[le][f];
                bc = fr.bc;
                flags = fr.flags;
                const int pt2 = (flags & 1) ? kl * N + (N - 1 - al) : pt;
                const double* m = fr.fm + pt;  // interface metric: 4 loads issued before anything depends on them
                fm[0] = m[0]; fm[1] = m[F]; fm[2] = m[2 * F]; fm[3] = m[3 * F];
                const double* t = fr.own + pt;
                const double* src = fr.nb + pt2;
#pragma unroll
                for (int v = 0; v < NTR; ++v) own[v] = t[v * F];
#pragma unroll
                for (int v = 0; v < NTR; ++v) nb[v] = src[v * F];
            }
            apply_xform<N>(a, flags, al, nb);
            double res[NFF];
            rusanov_point<N>(a, ei, d, side, pt, bc, own, nb, fm, res);
#pragma unroll
            for (int v = 0; v < NFF; ++v) *ffp(f * NFF + v, pt) = res[v];
        }
    }

    // ---- nodes ----
    const int ii = node % N, jj = (node / N) % N, kk = node / (N * N);
    double sg = 1.0, p = 1.0, gfac = 0.0, fo[3] = {0.0, 0.0, 0.0};
    long long qo = 0, vst = 0;
    double* const col = &sm.st[0][le][node];  // this node's column (rows RS apart)
    if (active) {
        const ElemInfo& ei = sm.ei[le];
        qo = ei.qoff + node;
        vst = ei.vstride;
        const long long mo = (long long)e * P + node;
        // the element's metric block [35][P] (Christoffel 0..26 | sqrtG | h 0..5 | 1/z_eta): one base, constant offsets
        const double* const mrow = a.chris + (long long)e * (NMETRIC * P) + node;
        double q[NVAR], h[6];
#pragma unroll
        for (int v = 0; v < NVAR; ++v) q[v] = qsrc[qo + v * vst];
        sg = mrow[R_SG * P];
#pragma unroll
        for (int c = 0; c < 6; ++c) h[c] = mrow[(R_H + c) * P];
        gfac = mrow[R_IDZ * P];
        double ch[27];  // all 27 Christoffel loads in flight together
        if (!WXB_CH3) {
#pragma unroll
            for (int c = 0; c < 27; ++c) ch[c] = mrow[c * P];
        }
        const double rho = q[0];
#if WXB_RCP
        const double irho = m_rcp(rho);
        const double u1 = q[1] * irho, u2 = q[2] * irho, w = q[3] * irho;
#else
        const double u1 = q[1] / rho, u2 = q[2] / rho, w = q[3] / rho;
#endif
        const double lg = m_log((WXB_RD / WXB_P0) * q[4], a.fasttab);
        p = WXB_P0 * m_exp(WXB_GAMMA * lg, a.fasttab);  // pde_euler_cubesphere.py:100
        // forcing (pde_euler_cubesphere.py:203-290)
        {
            const double t11 = rho * u1 * u1 + h[0] * p, t12 = rho * u1 * u2 + h[1] * p, t13 = rho * u1 * w + h[2] * p;
            const double t22 = rho * u2 * u2 + h[3] * p, t23 = rho * u2 * w + h[4] * p, t33 = rho * w * w + h[5] * p;
#pragma unroll
            for (int i = 0; i < 3; ++i) {
                double c[9];
#pragma unroll
                for (int m = 0; m < 9; ++m) c[m] = WXB_CH3 ? mrow[(9 * i + m) * P] : ch[9 * i + m];
                fo[i] = 2.0 * rho * (c[0] * u1 + c[1] * u2 + c[2] * w) + c[3] * t11 + 2.0 * c[4] * t12 + 2.0 * c[5] * t13 +
                        c[6] * t22 + 2.0 * c[7] * t23 + c[8] * t33;
            }
            if (a.has_damp) {
                const double dw = rho / a.tau0 * a.damp[mo];
                fo[0] += dw * (u1 - a.damp[EP + mo]);
                fo[1] += dw * (u2 - a.damp[2 * EP + mo]);
                fo[2] += dw * (w - a.damp[3 * EP + mo]);
            }
        }
        // pointwise fluxes of the three directions (pde_euler_cubesphere.py:72-124) -> own column
#pragma unroll
        for (int d = 0; d < 3; ++d) {
            const double ud = d == 0 ? u1 : (d == 1 ? u2 : w);
            const double hd0 = d == 0 ? h[0] : (d == 1 ? h[1] : h[2]);
            const double hd1 = d == 0 ? h[1] : (d == 1 ? h[3] : h[4]);
            const double hd2 = d == 0 ? h[2] : (d == 1 ? h[4] : h[5]);
            const double sgu = sg * ud;
            col[(d * 6 + 0) * RS] = sgu * q[0];
            col[(d * 6 + 1) * RS] = sgu * q[1] + sg * hd0 * p;
            col[(d * 6 + 2) * RS] = sgu * q[2] + sg * hd1 * p;
            col[(d * 6 + 3) * RS] = sgu * q[4];
            col[(d * 6 + 4) * RS] = sgu * q[3];
            col[(d * 6 + 5) * RS] = sg * hd2;
        }
#if WXB_FOLD_LOGP
        col[RLOGP * RS] = lg;  // log p = log p0 + gamma*lg (pde_euler_cubesphere.py:124); see O_DLP
#else
        col[RLOGP * RS] = WXB_LOG_P0 + WXB_GAMMA * lg;  // log p (pde_euler_cubesphere.py:124) without a second log
#endif
        col[RSGRHO * RS] = sg * rho;
    }
    __syncthreads();
    if (!active) return;

    // the epilogue's operands are requested now so that their latency hides behind the derivative
    double r0[NVAR], vv[NVAR];
    if (JVP) {
#pragma unroll
        for (int v = 0; v < NVAR; ++v) r0[v] = a.rhs0[qo + v * vst];
        if (a.mode == WXB_JVP_RAT) {
#pragma unroll
            for (int v = 0; v < NVAR; ++v) vv[v] = a.v[qo + v * vst];
        }
    }

    // ---- interior derivative (sum-factorised) + lift (correction) ----
    double acc[4] = {0.0, 0.0, 0.0, 0.0}, wacc = 0.0, grav = 0.0;
#pragma unroll
    for (int d = 0; d < 3; ++d) {
        const int idx = d == 0 ? ii : (d == 1 ? jj : kk);
        const int step = d == 0 ? 1 : (d == 1 ? N : N * N);
        const int pt = d == 0 ? kk * N + jj : (d == 1 ? kk * N + ii : jj * N + ii);
        const double* sfd = &sm.st[d * 6][le][node - idx * step];
        const double* slp = &sm.st[RLOGP][le][node - idx * step];
        double dv[7] = {0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int m = 0; m < N; ++m) {
            const double dm = s_op[C::O_DIFF + idx * N + m];
#pragma unroll
            for (int v = 0; v < 6; ++v) dv[v] += dm * sfd[v * RS + m * step];
            dv[6] += (WXB_FOLD_LOGP ? s_op[C::O_DLP + idx * N + m] : dm) * slp[m * step];
            if (d == 2) grav += s_op[C::O_HF + idx * N + m] * slp[RS + m * step];
        }
        const double cw = s_op[C::O_CORR + idx * 2], ce = s_op[C::O_CORR + idx * 2 + 1];
        // the reference adds the lift to the derivative field by field, then combines (rhs_dfr.py:106-139)
#pragma unroll
        for (int v = 0; v < NFF; ++v) dv[v] += cw * *ffp(2 * d * NFF + v, pt) + ce * *ffp((2 * d + 1) * NFF + v, pt);
        // d(log p) + lift of the own-side face log p = gamma * (D + c_w (x) ext_w + c_e (x) ext_e) lg + const(idx)
        if (WXB_FOLD_LOGP) dv[6] = WXB_GAMMA * dv[6] + s_op[C::O_KLP + idx];
#pragma unroll
        for (int v = 0; v < 4; ++v) acc[v] += dv[v];
        const double wpres = col[(d * 6 + 5) * RS];
        wacc += dv[4] + p * dv[5] + p * wpres * dv[6];
    }

    // ---- assemble ----
    const double inv_sg = m_rcp(sg);
    double r[NVAR];
    r[0] = -inv_sg * acc[0];
    r[1] = -inv_sg * acc[1] - fo[0];
    r[2] = -inv_sg * acc[2] - fo[1];
    r[3] = -inv_sg * wacc - (fo[2] + gfac * WXB_GRAVITY * inv_sg * grav);
    r[4] = -inv_sg * acc[3];
    // ---- epilogue: plain RHS, or the finite-difference Jacobian-vector product (matvec.py:57-88) ----
#pragma unroll
    for (int v = 0; v < NVAR; ++v) {
        const long long o = qo + v * vst;
        double x = r[v];
        if (JVP) {
            x = a.dt * (x - r0[v]) * a.inv_eps;  // matvec.py:66 divides by eps: same to one rounding
            if (a.mode == WXB_JVP_RAT) x = vv[v] - 0.5 * x;
        }
        a.out[o] = x;
    }
}

template <int N>
int launch_traces(const EvalArgs& a, cudaStream_t s) {
    const int ne = a.elist ? a.nelist : a.E;
    if (ne == 0) return 0;
    const int blocks = (ne + Cfg<N>::TEB - 1) / Cfg<N>::TEB;
    euler_traces<N><<<blocks, Cfg<N>::TTHREADS, 0, s>>>(a);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
template <int N>
int launch_main(const EvalArgs& a, cudaStream_t s) {
    const int ne = a.elist ? a.nelist : a.E;
    if (ne == 0) return 0;
    const int blocks = (ne + Cfg<N>::EB - 1) / Cfg<N>::EB;
    constexpr int smem = (int)sizeof(MainSmem<N>);
    if (a.mode >= 0)
        euler_main<N, true><<<blocks, Cfg<N>::THREADS, smem, s>>>(a);
    else
        euler_main<N, false><<<blocks, Cfg<N>::THREADS, smem, s>>>(a);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

template <int N>
int configure_main() {
    if (cudaFuncSetAttribute(euler_main<N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(MainSmem<N>)) != cudaSuccess) return -1;
    if (cudaFuncSetAttribute(euler_main<N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(MainSmem<N>)) != cudaSuccess) return -1;
    return 0;
}

}  // namespace

// opt the main kernel of order n into its dynamic shared memory size on the current device (once per handle)
int wxb_euler_configure(int n) {
    switch (n) {
        case 2: return configure_main<2>();
        case 3: return configure_main<3>();
        case 4: return configure_main<4>();
        case 5: return configure_main<5>();
        case 6: return configure_main<6>();
    }
    return -1;
}

// pack the trace records other handles need: sendbuf[rec][var][k][F]
__global__ void pack_kernel(const EvalArgs a, int nvar, int F, int nfaces, int elem_major) {
    const long long per_rec = (long long)nvar * a.nz * F;
    const long long total = per_rec * a.nsend;
    for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        const int rec = (int)(t / per_rec);
        int r = (int)(t % per_rec);
        int v, k;
        if (elem_major) {  // record [k][field][F], the order of the element-major trace array
            k = r / (nvar * F);
            r %= nvar * F;
            v = r / F;
        } else {           // record [field][k][F]
            v = r / (a.nz * F);
            r %= a.nz * F;
            k = r / F;
        }
        const int pt = r % F;
        const int cf = a.send[rec];
        const int c = cf >> 2, f = cf & 3;
        // column -> slab
        int s = 0;
        for (int u = 1; u < a.st.nslabs; ++u)
            if (c >= a.st.cbase[u]) s = u;
        const int cl = c - a.st.cbase[s];
        const int e = a.st.ebase[s] + k * a.st.ny[s] * a.nx + cl;
        a.sendbuf[t] = elem_major ? a.trace[(((long long)e * nfaces + f) * nvar + v) * F + pt]
                                  : a.trace[((long long)(v * nfaces + f) * a.E + e) * F + pt];
    }
}

int wxb_pack(const EvalArgs& a, int nvar, int F, int nfaces, int elem_major, cudaStream_t s) {
    if (a.nsend == 0) return 0;
    const long long total = (long long)a.nsend * nvar * a.nz * F;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 148 * 8) blocks = 148 * 8;
    pack_kernel<<<blocks, 256, 0, s>>>(a, nvar, F, nfaces, elem_major);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int wxb_euler_traces(const EvalArgs& a, int n, cudaStream_t s) {
    switch (n) {
        case 2: return launch_traces<2>(a, s);
        case 3: return launch_traces<3>(a, s);
        case 4: return launch_traces<4>(a, s);
        case 5: return launch_traces<5>(a, s);
        case 6: return launch_traces<6>(a, s);
    }
    return -1;
}
int wxb_euler_main(const EvalArgs& a, int n, cudaStream_t s) {
    switch (n) {
        case 2: return launch_main<2>(a, s);
        case 3: return launch_main<3>(a, s);
        case 4: return launch_main<4>(a, s);
        case 5: return launch_main<5>(a, s);
        case 6: return launch_main<6>(a, s);
    }
    return -1;
}
