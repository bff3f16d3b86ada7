// 3-D Euler DFR right-hand side on the cubed sphere -- the fused column-sweep kernel (sm_100a, FP64 CUDA-core FMA).
//
// One launch per evaluation and NO trace array.  A block owns a TJ x TI tile of element columns and walks up the
// column, one layer of TJ*TI elements per step, one thread per solution point of the layer.  Per step:
//
//   stage     the layer's own elements one level AHEAD (k+1) and the 2(TJ+TI) horizontal halo elements of level k are
//             read from the state (q, or q + eps*v for the FD Jacobian-vector product, wx_factory/solvers/matvec.py:57-66),
//             and their momenta, log(rho), log(rho theta) are left in shared memory
//             (solution_extrapolation's logs, wx_factory/rhs/rhs_dfr.py:50-71)
//   faces     every PHYSICAL face of the layer is evaluated once, by one thread per face point: both sides' traces
//             are extrapolated from the staged rows (own elements, halo elements -- through the pull table with the
//             panel-edge reversal / basis change of wx_factory/process_topology.py:128-175 -- or, on a multi-rank
//             handle, read from a ghost record), then the Rusanov flux (wx_factory/pde/fluxes.py:150-222,326-403,
//             507-582) with the rigid-lid mirror at the bottom and the top (wx_factory/rhs/rhs_dfr.py:265-268,
//             wx_factory/pde/pde_euler_cubesphere.py:152-156).  The face between level k and k+1 uses the staged
//             level k+1; its result is kept for the next step, so vertical faces are computed once, too.
//   nodes     pointwise fluxes and forcing (wx_factory/pde/pde_euler_cubesphere.py:72-124,203-290) as in euler3d.cu
//   assemble  sum-factorised derivative, lift, split rho-w form (wx_factory/rhs/rhs_dfr.py:25-34,89-139), epilogue
//
// Compared with the two-kernel path of euler3d.cu (traces kernel + main kernel): the state is read once from DRAM
// (the halo re-reads hit L2: neighbouring tiles sweep at the same time), no trace is ever written or read back
// (-67 B written, -80 B read per node at n = 4), interface fluxes are evaluated once per physical face inside the
// tile instead of twice, and -- because the block knows where it will be one step ahead -- the next level's metric
// block and state rows are requested into L2 a whole step before they are used, which decouples the DRAM latency
// from the per-thread register budget.  The price is the halo: 2(TJ+TI)/(TJ*TI) extra elements per element whose
// logs are recomputed by the neighbouring tile (with the table-driven log / exp of fastmath.cuh this is cheaper than
// what the trace kernel spent).
//
// Both sides of a tile-exterior face are computed by the two tiles that share it with the SAME device function on
// the SAME staged values, so they see bit-identical traces, as with the reference's single trace array.
#include "fastmath.cuh"
#include "wxb_internal.cuh"

// tile shape per order (elements per layer = threads / n^3) and resident blocks per SM the kernel is compiled for
#ifndef WXB_SW_TJ2
#define WXB_SW_TJ2 4
#endif
#ifndef WXB_SW_TI2
#define WXB_SW_TI2 4
#endif
#ifndef WXB_SW_TJ3
#define WXB_SW_TJ3 2
#endif
#ifndef WXB_SW_TI3
#define WXB_SW_TI3 4
#endif
#ifndef WXB_SW_TJ4
#define WXB_SW_TJ4 2
#endif
#ifndef WXB_SW_TI4
#define WXB_SW_TI4 2
#endif
#ifndef WXB_SW_TJ5
#define WXB_SW_TJ5 1
#endif
#ifndef WXB_SW_TI5
#define WXB_SW_TI5 2
#endif
#ifndef WXB_SW_TJ6
#define WXB_SW_TJ6 1
#endif
#ifndef WXB_SW_TI6
#define WXB_SW_TI6 2
#endif
#ifndef WXB_SW_MINB2
#define WXB_SW_MINB2 4
#endif
#ifndef WXB_SW_MINB3
#define WXB_SW_MINB3 2
#endif
#ifndef WXB_SW_MINB4
#define WXB_SW_MINB4 2
#endif
#ifndef WXB_SW_MINB5
#define WXB_SW_MINB5 2
#endif
#ifndef WXB_SW_MINB6
#define WXB_SW_MINB6 1
#endif
#ifndef WXB_SW_ALIAS
#define WXB_SW_ALIAS 0  // 1: the flux fields share their shared memory with the halo rows and the spent own rows (two more
                        // barriers per step, 30 KB less per block at n = 4: 3 resident blocks instead of 2)
#endif
#ifndef WXB_SW_SPLITORDER
#define WXB_SW_SPLITORDER 1  // half of the block's warps run the node pass before the face pass, the other half after it:
                             // the node pass waits for 40 loads per thread, the face pass is arithmetic on shared memory, so
                             // each half's loads fly while the other half computes (needs WXB_SW_ALIAS = 0)
#endif
#ifndef WXB_SW_JVP_EARLY_OWN
#define WXB_SW_JVP_EARLY_OWN 0  // 1: JVP mode requests the own rows (q and v) of the next level ahead of the assemble phase too
                                // (measured slower: JVP 1.41 vs 1.37 ms, the 20 registers held cost more than the latency hidden)
#endif
#ifndef WXB_SW_CH3
#define WXB_SW_CH3 0  // 1: Christoffel symbols loaded 9 at a time where they are used (fewer registers, less memory parallelism)
#endif
#ifndef WXB_SW_SKIP
#define WXB_SW_SKIP 0  // TIMING EXPERIMENTS ONLY (compile-time, never set in a shipped build; results are wrong): bit 0 skip
                       // the face pass, 1 the node pass, 2 the assemble phase's arithmetic, 3 the halo staging
#endif
#ifndef WXB_SW_PREFETCH
#define WXB_SW_PREFETCH 3  // request the metric block / state rows of a later level into L2: 0 off, 1 one prefetch.global.L2
                           // per 128 bytes, 2 one per 32-byte sector, 3 one bulk prefetch per element (TMA engine)
#endif
#ifndef WXB_SW_PFDIST
#define WXB_SW_PFDIST 1    // how many levels ahead
#endif

namespace {

constexpr int NVAR = 5;
constexpr int NTR = 6;  // trace fields per face point: rho, u1, u2, w, rho theta, sound speed

template <int N>
struct SCfg {
    static constexpr int P = N * N * N, F = N * N;
    static constexpr int TJ = N == 2 ? WXB_SW_TJ2 : N == 3 ? WXB_SW_TJ3 : N == 4 ? WXB_SW_TJ4 : N == 5 ? WXB_SW_TJ5 : WXB_SW_TJ6;
    static constexpr int TI = N == 2 ? WXB_SW_TI2 : N == 3 ? WXB_SW_TI3 : N == 4 ? WXB_SW_TI4 : N == 5 ? WXB_SW_TI5 : WXB_SW_TI6;
    static constexpr int MINB = N == 2 ? WXB_SW_MINB2 : N == 3 ? WXB_SW_MINB3 : N == 4 ? WXB_SW_MINB4 : N == 5 ? WXB_SW_MINB5 : WXB_SW_MINB6;
    static constexpr int NE = TJ * TI;         // own elements per layer
    static constexpr int NH = 2 * (TJ + TI);   // halo elements per layer: W[TJ] E[TJ] S[TI] N[TI]
    static constexpr int THREADS = NE * P;
    static constexpr int NXF = TJ * (TI + 1);  // physical x1-faces of a layer, index a*(TI+1) + pos
    static constexpr int NYF = (TJ + 1) * TI;  // physical x2-faces, index pos*TI + b
    static constexpr int NFACE = NXF + NYF + NE;  // + the face above every own element
    // offsets into the operator block (wxb_api.cu: build)
    static constexpr int O_DIFF = 2 * N, O_CORR = 2 * N + N * N, O_HF = 4 * N + N * N;
    static constexpr int O_DLP = 4 * N + 2 * N * N, O_KLP = O_DLP + N * N;
    static constexpr int NOPS = 5 * N + 3 * N * N;
};

// where a halo element's values come from
struct HaloDesc {
    long long off0;     // kind 1: (var 0, level 0, node 0) of the neighbour element in the state; kind 2: field 0, level 0,
                        // point 0 of the ghost record in the trace array
    long long vstride;  // kind 1: distance between variables in the neighbour's slab
    long long lvl;      // distance between levels (doubles)
    int flags;          // bit0 = reversed along the edge, bits 8.. = 1 + xform row (partition.py)
    short fq;           // kind 1: the neighbour's face that touches us (0 W, 1 E, 2 S, 3 N)
    short kind;         // 0 none, 1 staged from the state, 2 ghost record
};

struct FaceDesc {
    const double* fm;   // interface metric (4 components x F) of this face at level 0 (vertical faces: the face ABOVE level 0)
    long long fmk;      // distance between levels
    short lo, hi;       // min-side / max-side element: own slot 0..NE-1, 100 + halo slot, -1 = face does not exist
};

constexpr int R_M = 0, R_LRHO = 3, R_LRT = 4, NSTAGE = 5;  // staged rows: rho u1, rho u2, rho w, log rho, log(rho theta)
// own elements also keep rho and rho theta (of q, or of q + eps v): the node pass of the level then reads its state from
// shared memory instead of loading q (and v) a second time (not with WXB_SW_ALIAS, where the flux fields overwrite the rows)
// (n = 5: the two extra rows would push the block over half an SM's shared memory)
template <int N>
struct QSmem {
    static constexpr bool value = !WXB_SW_ALIAS && N != 5;
};
constexpr int R_RHO = 5, R_RT = 6;
constexpr int NFLUX = 20;  // per direction F_rho F_u1 F_u2 F_theta w_adv w_pres; then log(Rd/p0 rho theta), sqrtG rho

// staged rows and flux fields in one array (doubles):
//   so(b)  [NSTAGE][NE][P]  own elements of the levels of parity b   (written one step ahead, dead after `faces`)
//   sh     [NSTAGE][NH][P]  halo elements of the current level       (dead after `faces`)
//   fl(k)  [NFLUX][NE][P]   flux fields of the current level         (written by `nodes`, read by `assemble`)
// WXB_SW_ALIAS: so(0) | sh | pad | so(1), and fl(k) starts at so(0) for even k, at sh for odd k: it overwrites the spent
// own rows and the halo rows but never the rows staged for the next level.
template <int N>
struct SweepLayout {
    using C = SCfg<N>;
    static constexpr int SO = (QSmem<N>::value ? 7 : 5) * C::NE * C::P, SH = NSTAGE * C::NH * C::P, FL = NFLUX * C::NE * C::P;
    static constexpr int PAD = FL > SO + SH ? FL - SO - SH : 0;
    static constexpr int TOTAL = WXB_SW_ALIAS ? 2 * SO + SH + PAD : 2 * SO + SH + FL;
    __device__ static constexpr int so(int b) { return WXB_SW_ALIAS ? (b ? SO + SH + PAD : 0) : b * SO; }
    __device__ static constexpr int sh() { return WXB_SW_ALIAS ? SO : 2 * SO; }
    __device__ static constexpr int fl(int k) { return WXB_SW_ALIAS ? ((k & 1) ? SO : 0) : 2 * SO + SH; }
};

template <int N>
struct SweepSmem {
    using C = SCfg<N>;
    alignas(16) double buf[SweepLayout<N>::TOTAL];
    double ffx[C::NXF][7][C::F];        // interface results per physical face: 5 common values, w_pres* / p of the
    double ffy[C::NYF][7][C::F];        //   min-side (L) element [5] and of the max-side (R) element [6]
    double ffu[C::NE][7][C::F];         // face above each own element
    double ffd[C::NE][6][C::F];         // face below: last step's ffu with [5] = its [6]
    double op[C::NOPS];
    double tab[WXB_FASTTAB_DOUBLES];
    HaloDesc hd[C::NH];
    FaceDesc fd[C::NFACE];
};

// Staged element rows are stored SWIZZLED: node (k, j, i) of an element sits at k n^2 + ((j + k) mod n) n + i.  With the
// plain order the 16 lines an x2-face reads (one per (k, i), n = 4) start 16 doubles = 32 banks apart for successive k
// and collide 4-way in shared memory; rotating j by k spreads them over all banks, and the lines along x1 stay
// contiguous (128-bit loads) and those along x3 conflict-free.
template <int N>
__device__ __forceinline__ int swz(int k, int j, int i) {
    int t = j + k;
    t -= t >= N ? N : 0;
    return k * N * N + t * N + i;
}
template <int N>
__device__ __forceinline__ int swz_node(int node) {
    const int k = node / (N * N), r = node - k * N * N, j = r / N;
    return swz<N>(k, j, r - j * N);
}

// one trace point from the staged rows of an element (rows `rs` apart): extrapolation of the momenta and of the two
// logs along the line of face point pt = c1*n + c0 of direction d (x1-face pt = (k,j), x2-face (k,i), x3-face (j,i);
// operators.py:161-165), then the fields the flux consumes: rho, u1, u2, w, rho theta and the sound speed
// a = sqrt(gamma p / rho) = sqrt(gamma p0) exp((gamma (log(rho theta) + log(Rd/p0)) - log rho) / 2)
template <int N>
__device__ __forceinline__ void form_trace(const double* __restrict__ mp, int rs, const double* __restrict__ ext, int d, int pt,
                                           const double* __restrict__ etab, double out[NTR]) {
    const int c1 = pt / N, c0 = pt - c1 * N;
    double a1 = 0.0, a2 = 0.0, a3 = 0.0, l0 = 0.0, l4 = 0.0;
    if (N % 2 == 0 && d == 0) {
        // a line along x1 is N consecutive doubles (16-byte aligned for even N): 128-bit loads; same fma order as below
        const double2* x2 = reinterpret_cast<const double2*>(mp + swz<N>(c1, c0, 0));
        double xv[5][N];
#pragma unroll
        for (int r = 0; r < 5; ++r)
#pragma unroll
            for (int m = 0; m < N / 2; ++m) {
                const double2 t = x2[(r * rs) / 2 + m];
                xv[r][2 * m] = t.x;
                xv[r][2 * m + 1] = t.y;
            }
#pragma unroll
        for (int m = 0; m < N; ++m) {
            const double w = ext[m];
            a1 = fma(w, xv[0][m], a1);
            a2 = fma(w, xv[1][m], a2);
            a3 = fma(w, xv[2][m], a3);
            l0 = fma(w, xv[3][m], l0);
            l4 = fma(w, xv[4][m], l4);
        }
    } else {
#pragma unroll
        for (int m = 0; m < N; ++m) {
            const double w = ext[m];
            const double* x = mp + (d == 0 ? swz<N>(c1, c0, m) : (d == 1 ? swz<N>(c1, m, c0) : swz<N>(m, c1, c0)));
            a1 = fma(w, x[0], a1);
            a2 = fma(w, x[rs], a2);
            a3 = fma(w, x[2 * rs], a3);
            l0 = fma(w, x[R_LRHO * rs], l0);
            l4 = fma(w, x[R_LRT * rs], l4);
        }
    }
    const double rho = wxbfm::fast_exp(l0, etab);
    const double r = wxbfm::fast_rcp(rho);
    out[0] = rho;
    out[1] = a1 * r;
    out[2] = a2 * r;
    out[3] = a3 * r;
    out[4] = wxbfm::fast_exp(l4, etab);
    out[5] = WXB_SQRT_GAMMA_P0 * wxbfm::fast_exp(0.5 * (WXB_GAMMA * (l4 + WXB_LOG_RD_P0) - l0), etab);
}

// Rusanov flux of one physical face point from its min-side (L) and max-side (R) traces (fluxes.py:150-222,326-403,
// 507-582): c[0..3] = F* rows rho, rho u1, rho u2, rho theta; c[4] = wflux_adv*; w5L / w5R = wflux_pres* / p of the
// L / R element (the reference keeps a different value on each side, fluxes.py:402-403).  mirL / mirR: that side is
// the mirrored ghost of the other one and its normal velocity is reflected (pde_euler_cubesphere.py:152-156).
__device__ __forceinline__ void rusanov_lr(int d, const double L[NTR], const double R[NTR], bool mirL, bool mirR, const double fm[4],
                                           double c[5], double& w5L, double& w5R) {
    const double sg = fm[0], hd0 = fm[1], hd1 = fm[2], hd2 = fm[3];
    const double hdd = d == 0 ? hd0 : (d == 1 ? hd1 : hd2);
    double uL = d == 0 ? L[1] : (d == 1 ? L[2] : L[3]);
    double uR = d == 0 ? R[1] : (d == 1 ? R[2] : R[3]);
    if (mirL) uL = -uR;
    if (mirR) uR = -uL;
    constexpr double inv_gamma = 1.0 / WXB_GAMMA;
    const double pL = L[5] * L[5] * (L[0] * inv_gamma), pR = R[5] * R[5] * (R[0] * inv_gamma);  // a^2 rho / gamma
    const double sh = wxbfm::fast_sqrt(hdd);
    const double eig = fmax(fabs(uL) + sh * L[5], fabs(uR) + sh * R[5]);  // |u| + sqrt(h^dd gamma p / rho)
    const double sguL = sg * uL, sguR = sg * uR, es = eig * sg;
    const double mL1 = L[0] * L[1], mL2 = L[0] * L[2], mL3 = L[0] * L[3];
    const double mR1 = R[0] * R[1], mR2 = R[0] * R[2], mR3 = R[0] * R[3];
    c[0] = 0.5 * (sguL * L[0] + sguR * R[0] - es * (R[0] - L[0]));
    c[1] = 0.5 * ((sguL * mL1 + sg * hd0 * pL) + (sguR * mR1 + sg * hd0 * pR) - es * (mR1 - mL1));
    c[2] = 0.5 * ((sguL * mL2 + sg * hd1 * pL) + (sguR * mR2 + sg * hd1 * pR) - es * (mR2 - mL2));
    c[3] = 0.5 * (sguL * L[4] + sguR * R[4] - es * (R[4] - L[4]));
    c[4] = 0.5 * (sguL * mL3 + sguR * mR3 - es * (mR3 - mL3));
    const double num = 0.5 * (sg * hd2 * pL + sg * hd2 * pR);
    w5L = num * wxbfm::fast_rcp(pL);
    w5R = num * wxbfm::fast_rcp(pR);
}

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
// one instruction for a whole contiguous block (bytes: a multiple of 16, address 16-byte aligned)
__device__ __forceinline__ void prefetch_bulk_l2(const void* p, unsigned bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

// ------------------------------------------------------------------------------------------------------
template <int N, bool JVP>
__global__ void __launch_bounds__(SCfg<N>::THREADS, SCfg<N>::MINB) euler_sweep(const EvalArgs a) {
    using C = SCfg<N>;
    constexpr int P = C::P, F = C::F, TJ = C::TJ, TI = C::TI, NE = C::NE, NH = C::NH, NT = C::THREADS;
    constexpr int RSO = NE * P, RSH = NH * P;  // distance between staged rows (own / halo)
    extern __shared__ __align__(16) unsigned char smem_raw[];
    SweepSmem<N>& sm = *reinterpret_cast<SweepSmem<N>*>(smem_raw);
    using LY = SweepLayout<N>;
    double* const sbuf = sm.buf;
    const int tid = threadIdx.x, le = tid / P, node = tid - le * P;
    const double* const s_op = sm.op;
    const double* const ltab = sm.tab;
    const double* const etab = sm.tab + WXB_EXPTAB_OFF;

    // ---- once per tile -------------------------------------------------------------------------------
    for (int t = tid; t < C::NOPS; t += NT) sm.op[t] = a.ops[t];
    for (int t = tid; t < WXB_FASTTAB_DOUBLES; t += NT) sm.tab[t] = a.fasttab[t];
    const int tile = a.tlist ? a.tlist[blockIdx.x] : (int)blockIdx.x;
    const int nx = a.nx, nz = a.nz;
    const int nti = (nx + TI - 1) / TI;
    int s = 0, trem = tile;
    for (int u = 0; u + 1 < a.st.nslabs; ++u) {
        const int nt = ((a.st.ny[u] + TJ - 1) / TJ) * nti;
        if (s == u && trem >= nt) {
            trem -= nt;
            s = u + 1;
        }
    }
    const int ny = a.st.ny[s];
    const int tj0 = (trem / nti) * TJ, ti0 = (trem % nti) * TI;
    const int aj = min(TJ, ny - tj0), ai = min(TI, nx - ti0);  // rows / columns of the tile that exist
    const long long lvl = (long long)ny * nx * P;              // doubles per level of one variable in this slab
    const long long vstride = (long long)nz * lvl;
    const int sa = le / TI, sb = le - sa * TI;
    const bool active = sa < aj && sb < ai;
    const long long col_e = (long long)(tj0 + sa) * nx + ti0 + sb;        // element within level 0 of the slab
    const long long e0 = a.st.ebase[s] + col_e;                            // ... and in the handle's element order
    const long long qoff0 = (long long)NVAR * P * a.st.ebase[s] + col_e * P + node;

    if (tid < NH) {
        const int h = tid;
        int face, ha, hb;
        bool ok;
        if (h < TJ) { face = 0; ha = h; hb = 0; ok = ha < aj; }
        else if (h < 2 * TJ) { face = 1; ha = h - TJ; hb = ai - 1; ok = ha < aj; }
        else if (h < 2 * TJ + TI) { face = 2; ha = 0; hb = h - 2 * TJ; ok = hb < ai; }
        else { face = 3; ha = aj - 1; hb = h - 2 * TJ - TI; ok = hb < ai; }
        HaloDesc d;
        d.off0 = d.vstride = d.lvl = 0;
        d.flags = 0;
        d.fq = 0;
        d.kind = 0;
        if (ok) {
            const int c = a.st.cbase[s] + (tj0 + ha) * nx + ti0 + hb;
            const int4 t = a.nbr[(long long)face * a.C + c];  // base, level stride, field stride, flags (rows of F doubles)
            d.flags = t.w;
            if ((long long)t.x < (long long)NTR * 6 * a.E) {  // a local element: [E][6 faces][6 fields] numbering
                const int en = t.x / (6 * NTR);
                d.fq = (short)((t.x / NTR) % 6);
                const ElemPos ep = decode_elem(a.st, en, nz, nx, P, NVAR);
                d.off0 = ep.qoff;
                d.vstride = ep.vstride;
                d.lvl = (long long)(t.y / (6 * NTR)) * P;
                d.kind = 1;
            } else {  // a ghost record [rec][k][field][F] received from another rank
                d.off0 = (long long)t.x * F;
                d.lvl = (long long)t.y * F;
                d.kind = 2;
            }
        }
        sm.hd[h] = d;
    }
    if (tid < C::NFACE) {
        const int f = tid;
        FaceDesc d;
        d.fm = nullptr;
        d.fmk = 0;
        d.lo = d.hi = -1;
        if (f < C::NXF) {
            const int fa = f / (TI + 1), pos = f - fa * (TI + 1);
            if (fa < aj && pos <= ai) {
                d.lo = pos > 0 ? fa * TI + pos - 1 : 100 + fa;
                d.hi = pos < ai ? fa * TI + pos : 100 + TJ + fa;
                d.fm = a.fmi + a.st.fmi[s] + ((long long)(tj0 + fa) * (nx + 1) + ti0 + pos) * (4 * F);
                d.fmk = (long long)ny * (nx + 1) * (4 * F);
            }
        } else if (f < C::NXF + C::NYF) {
            const int g = f - C::NXF, pos = g / TI, fb = g - pos * TI;
            if (fb < ai && pos <= aj) {
                d.lo = pos > 0 ? (pos - 1) * TI + fb : 100 + 2 * TJ + fb;
                d.hi = pos < aj ? pos * TI + fb : 100 + 2 * TJ + TI + fb;
                d.fm = a.fmj + a.st.fmj[s] + ((long long)(tj0 + pos) * nx + ti0 + fb) * (4 * F);
                d.fmk = (long long)(ny + 1) * nx * (4 * F);
            }
        } else {
            const int slot = f - C::NXF - C::NYF, fa = slot / TI, fb = slot - fa * TI;
            if (fa < aj && fb < ai) {
                d.lo = d.hi = slot;
                d.fmk = (long long)ny * nx * (4 * F);
                d.fm = a.fmk + a.st.fmk[s] + d.fmk + ((long long)(tj0 + fa) * nx + ti0 + fb) * (4 * F);  // face above level 0
            }
        }
        sm.fd[f] = d;
    }

    // ---- stage: state rows -> momenta and logs in shared memory -----------------------------------------
    // In two halves: the LOADS of the next step's rows (own elements of level ko, halo elements of level kh; -1 = none)
    // are issued before the assemble phase of the current step, which hides their latency; logs and shared-memory
    // stores follow after it.  HR = halo (element, node) pairs per thread.
    constexpr int HR = (NH * P + NT - 1) / NT;
    // JVP mode stages q + eps*v: twice the operands to hold, so there the rows are requested where they are used
    // (register budget: 128 at two resident blocks)
    constexpr bool EARLY = !JVP && HR <= 2;      // halo rows requested early
    constexpr bool EARLY_OWN = WXB_SW_JVP_EARLY_OWN ? HR <= 2 : EARLY;  // own rows requested early
    double xo[NVAR], xh[HR][NVAR];
    double vo[JVP ? NVAR : 1], vh[HR][JVP ? NVAR : 1];
    auto halo_load1 = [&](int kh, int r) {
        if (WXB_SW_SKIP & 8) return;
        const int t = tid + r * NT;
        const int h = t / P, hn = t - h * P;
        if (t < NH * P && sm.hd[h].kind == 1) {
            const HaloDesc& hd = sm.hd[h];
            const long long o = hd.off0 + (long long)kh * hd.lvl + hn;
#pragma unroll
            for (int v = 0; v < NVAR; ++v) xh[r][v] = a.q[o + v * hd.vstride];
            if (JVP) {
#pragma unroll
                for (int v = 0; v < NVAR; ++v) vh[r][v] = a.v[o + v * hd.vstride];
            }
        }
    };
    auto halo_load = [&](int kh) {
#pragma unroll
        for (int r = 0; r < HR; ++r) halo_load1(kh, r);
    };
    auto own_load = [&](int ko) {
        if (active && ko >= 0) {
            const long long o = qoff0 + (long long)ko * lvl;
#pragma unroll
            for (int v = 0; v < NVAR; ++v) xo[v] = a.q[o + v * vstride];
            if (JVP) {
#pragma unroll
                for (int v = 0; v < NVAR; ++v) vo[v] = a.v[o + v * vstride];
            }
        }
    };
    auto stage_load = [&](int ko, int kh) {
        if (EARLY_OWN) own_load(ko);
        if (EARLY && kh >= 0) halo_load(kh);
    };
    auto stage_put = [&](double* dst, int rs, double x[NVAR], const double* dv, bool own) {
        if (JVP) {
#pragma unroll
            for (int v = 0; v < NVAR; ++v) x[v] += a.eps * dv[v];
        }
        if (QSmem<N>::value && own) {
            dst[R_RHO * rs] = x[0];
            dst[R_RT * rs] = x[4];
        }
        dst[0] = x[1];
        dst[rs] = x[2];
        dst[2 * rs] = x[3];
        dst[R_LRHO * rs] = wxbfm::fast_log(x[0], ltab);
        dst[R_LRT * rs] = wxbfm::fast_log(x[4], ltab);
    };
    auto stage_store = [&](int ko, int kh) {
        if (!EARLY_OWN) own_load(ko);
        if (active && ko >= 0) stage_put(sbuf + LY::so(ko & 1) + le * P + swz_node<N>(node), RSO, xo, vo, true);
        if (kh >= 0) {
#pragma unroll
            for (int r = 0; r < HR; ++r) {
                if (!EARLY) halo_load1(kh, r);
                const int t = tid + r * NT;
                const int h = t / P, hn = t - h * P;
                if (t < NH * P && sm.hd[h].kind == 1 && !(WXB_SW_SKIP & 8)) stage_put(sbuf + LY::sh() + h * P + swz_node<N>(hn), RSH, xh[r], vh[r], false);
            }
        }
    };

    // ---- faces of the layer whose lower level is k (zonly: only the faces above it) -------------------------
    auto faces = [&](int k, bool zonly) {
        for (int fp = tid; fp < C::NFACE * F; fp += NT) {
            const int f = fp / F, pt = fp - f * F;
            const int d = f < C::NXF ? 0 : (f < C::NXF + C::NYF ? 1 : 2);
            if (zonly && d != 2) continue;
            const FaceDesc& fd = sm.fd[f];
            const int lo = fd.lo, hi = fd.hi;
            if (lo < 0) continue;
            const double* fmp = fd.fm + (long long)k * fd.fmk + pt;
            double fm[4];
#pragma unroll
            for (int c = 0; c < 4; ++c) fm[c] = fmp[c * F];
            const int kl = pt / N, al = pt - kl * N;
            double L[NTR], R[NTR];
            bool mirL = false, mirR = false;
            // one side: where its staged rows are, which line, which end
            auto side = [&](int code, int own_side, double out[NTR]) {
                const double* mp;
                const double* ext;
                int rs, dl, ptl, flags = 0;
                const double* ghost = nullptr;
                if (code < 100) {
                    mp = sbuf + LY::so(k & 1) + code * P;
                    rs = RSO;
                    ext = s_op + own_side * N;
                    dl = d;
                    ptl = pt;
                } else {
                    const HaloDesc& hd = sm.hd[code - 100];
                    flags = hd.flags;
                    ptl = (flags & 1) ? kl * N + (N - 1 - al) : pt;
                    mp = sbuf + LY::sh() + (code - 100) * P;
                    rs = RSH;
                    ext = s_op + (hd.fq & 1) * N;
                    dl = hd.fq >> 1;
                    if (hd.kind == 2) ghost = a.trace + hd.off0 + (long long)k * hd.lvl + ptl;
                }
                if (ghost) {
#pragma unroll
                    for (int v = 0; v < NTR; ++v) out[v] = ghost[v * F];
                } else {
                    form_trace<N>(mp, rs, ext, dl, ptl, etab, out);
                }
                const int xid = flags >> 8;
                if (xid) {  // the neighbour's (u1, u2) in this panel's basis (process_topology.py:138-175)
                    const double* m = a.xform + ((long long)(xid - 1) * N + al) * 4;
                    const double b1 = m[0] * out[1] + m[1] * out[2];
                    const double b2 = m[2] * out[1] + m[3] * out[2];
                    out[1] = b1;
                    out[2] = b2;
                }
            };
            if (d == 2) {
                const bool haveL = k >= 0, haveR = k + 1 < nz;
                if (haveL) form_trace<N>(sbuf + LY::so(k & 1) + lo * P, RSO, s_op + N, 2, pt, etab, L);
                if (haveR) form_trace<N>(sbuf + LY::so((k + 1) & 1) + lo * P, RSO, s_op, 2, pt, etab, R);
                if (!haveL) {  // rigid bottom: the ghost mirrors the real face (rhs_dfr.py:265-268)
#pragma unroll
                    for (int v = 0; v < NTR; ++v) L[v] = R[v];
                    mirL = true;
                }
                if (!haveR) {
#pragma unroll
                    for (int v = 0; v < NTR; ++v) R[v] = L[v];
                    mirR = true;
                }
            } else {
                side(lo, 1, L);
                side(hi, 0, R);
            }
            double c[5], w5L, w5R;
            rusanov_lr(d, L, R, mirL, mirR, fm, c, w5L, w5R);
            double* dst = d == 0 ? &sm.ffx[f][0][pt] : (d == 1 ? &sm.ffy[f - C::NXF][0][pt] : &sm.ffu[f - C::NXF - C::NYF][0][pt]);
            if (d == 2) {  // what was the face above becomes the face below
                double* dd = &sm.ffd[f - C::NXF - C::NYF][0][pt];
#pragma unroll
                for (int v = 0; v < 5; ++v) dd[v * F] = dst[v * F];
                dd[5 * F] = dst[6 * F];
            }
#pragma unroll
            for (int v = 0; v < 5; ++v) dst[v * F] = c[v];
            dst[5 * F] = w5L;
            dst[6 * F] = w5R;
        }
    };

    const int k0 = a.k0, k1 = a.k1;
    __syncthreads();  // descriptors, operators, tables
    if (k0 > 0) {
        stage_load(k0 - 1, -1);
        stage_store(k0 - 1, -1);
    }
    stage_load(k0, -1);
    stage_store(k0, -1);
    stage_load(k0 + 1 < nz ? k0 + 1 : -1, k0);  // the first step's rows: in flight during the face pass below
    __syncthreads();
    faces(k0 - 1, true);  // the face below the first level (rigid bottom, or the level below a chunk)
    __syncthreads();

    const int ii = node % N, jj = (node / N) % N, kk = node / (N * N);
    const long long EP = (long long)a.E * P;
    constexpr int RS = NE * P;

    for (int k = k0; k < k1; ++k) {
        stage_store(k + 1 < nz ? k + 1 : -1, k);
#if WXB_SW_PREFETCH
        if (active && k + WXB_SW_PFDIST < k1) {  // a later step's metric block (35 rows of this element, contiguous) into L2
            const char* mb = (const char*)(a.chris + (e0 + (long long)(k + WXB_SW_PFDIST) * ny * nx) * (35 * P));
            constexpr int PFS = WXB_SW_PREFETCH == 2 ? 32 : 128;
            if (WXB_SW_PREFETCH == 3 && (35 * P * 8) % 16 == 0) {  // bulk form: 16-byte granularity (even n)
                if (node == 0) prefetch_bulk_l2(mb, 35 * P * 8);
            } else {
                for (int l = node * PFS; l < 35 * P * 8; l += P * PFS) prefetch_l2(mb + l);
            }
            if (k + 1 + WXB_SW_PFDIST < nz && node < NVAR) {  // the state rows staged by that step
                const char* qb = (const char*)(a.q + qoff0 - node + (long long)(k + 1 + WXB_SW_PFDIST) * lvl + node * vstride);
                for (int l = 0; l < P * 8; l += PFS) prefetch_l2(qb + l);
                if (JVP) {
                    const char* vb = (const char*)(a.v + qoff0 - node + (long long)(k + 1 + WXB_SW_PFDIST) * lvl + node * vstride);
                    for (int l = 0; l < P * 8; l += PFS) prefetch_l2(vb + l);
                }
            }
        }
#endif
        __syncthreads();
        constexpr bool SPLIT = WXB_SW_SPLITORDER && !WXB_SW_ALIAS;
        const bool node_first = SPLIT && ((tid >> 5) & 4);  // warps 4..7 of every 8: one of each kind per scheduler
        if (!SPLIT) {
            if (!(WXB_SW_SKIP & 1)) faces(k, false);
            if (WXB_SW_ALIAS) __syncthreads();  // the flux fields overwrite rows the face pass reads
        }
        double* const col = sbuf + LY::fl(k) + le * P + node;

        // ---- nodes: forcing and the flux fields of the three directions -> own column of fl ----------------
        double sg = 1.0, p = 1.0, gfac = 0.0, fo[3] = {0.0, 0.0, 0.0};
        const long long qo = qoff0 + (long long)k * lvl;
        auto nodes = [&]() {
        if (active && !(WXB_SW_SKIP & 2)) {
            const long long e = e0 + (long long)k * ny * nx;
            const double* const mrow = a.chris + e * (35 * P) + node;  // [35][P]: Christoffel 0..26 | sqrtG | h 0..5 | 1/z_eta
            double q[NVAR], h[6], ch[WXB_SW_CH3 ? 1 : 27];
            if (QSmem<N>::value) {  // staged one step ahead (q, or q + eps v)
                const double* sq = sbuf + LY::so(k & 1) + le * P + swz_node<N>(node);
                q[0] = sq[R_RHO * RSO];
                q[1] = sq[0];
                q[2] = sq[RSO];
                q[3] = sq[2 * RSO];
                q[4] = sq[R_RT * RSO];
            } else {
#pragma unroll
                for (int v = 0; v < NVAR; ++v) q[v] = a.q[qo + v * vstride];
                if (JVP) {
                    double dv[NVAR];
#pragma unroll
                    for (int v = 0; v < NVAR; ++v) dv[v] = a.v[qo + v * vstride];
#pragma unroll
                    for (int v = 0; v < NVAR; ++v) q[v] += a.eps * dv[v];
                }
            }
            sg = mrow[27 * P];
#pragma unroll
            for (int c = 0; c < 6; ++c) h[c] = mrow[(28 + c) * P];
            gfac = mrow[34 * P];
            if (!WXB_SW_CH3) {
#pragma unroll
                for (int c = 0; c < 27; ++c) ch[c] = mrow[c * P];
            }
            const double rho = q[0];
            const double irho = wxbfm::fast_rcp(rho);
            const double u1 = q[1] * irho, u2 = q[2] * irho, w = q[3] * irho;
            const double lg = wxbfm::fast_log((WXB_RD / WXB_P0) * q[4], ltab);
            p = WXB_P0 * wxbfm::fast_exp(WXB_GAMMA * lg, etab);  // pde_euler_cubesphere.py:100
            {
                const double t11 = rho * u1 * u1 + h[0] * p, t12 = rho * u1 * u2 + h[1] * p, t13 = rho * u1 * w + h[2] * p;
                const double t22 = rho * u2 * u2 + h[3] * p, t23 = rho * u2 * w + h[4] * p, t33 = rho * w * w + h[5] * p;
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    double c[9];
#pragma unroll
                    for (int m = 0; m < 9; ++m) c[m] = WXB_SW_CH3 ? mrow[(9 * i + m) * P] : ch[WXB_SW_CH3 ? 0 : 9 * i + m];
                    fo[i] = 2.0 * rho * (c[0] * u1 + c[1] * u2 + c[2] * w) + c[3] * t11 + 2.0 * c[4] * t12 + 2.0 * c[5] * t13 +
                            c[6] * t22 + 2.0 * c[7] * t23 + c[8] * t33;
                }
                if (a.has_damp) {  // Rayleigh damping of DCMIP 2-x (init/dcmip.py:676-755)
                    const long long mo = e * P + node;
                    const double dw = rho / a.tau0 * a.damp[mo];
                    fo[0] += dw * (u1 - a.damp[EP + mo]);
                    fo[1] += dw * (u2 - a.damp[2 * EP + mo]);
                    fo[2] += dw * (w - a.damp[3 * EP + mo]);
                }
            }
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
            col[18 * RS] = lg;
            col[19 * RS] = sg * rho;
        }
        };
        if (SPLIT) {
            if (node_first) nodes();
            if (!(WXB_SW_SKIP & 1)) faces(k, false);
            if (!node_first) nodes();
        } else {
            nodes();
        }
        __syncthreads();
        // the next step's rows (own level k+2, halo level k+1): requested now, used after the assemble phase
        if (k + 1 < k1) stage_load(k + 2 < nz ? k + 2 : -1, k + 1);

        if (active) {
            double r0[NVAR], vv[NVAR];
            if (JVP) {
#pragma unroll
                for (int v = 0; v < NVAR; ++v) r0[v] = a.rhs0[qo + v * vstride];
                if (a.mode == WXB_JVP_RAT) {
#pragma unroll
                    for (int v = 0; v < NVAR; ++v) vv[v] = a.v[qo + v * vstride];
                }
            }
            // ---- interior derivative (sum-factorised) + lift ----
            double acc[4] = {0.0, 0.0, 0.0, 0.0}, wacc = 0.0, grav = 0.0;
#pragma unroll
            for (int d = 0; d < ((WXB_SW_SKIP & 4) ? 0 : 3); ++d) {
                const int idx = d == 0 ? ii : (d == 1 ? jj : kk);
                const int step = d == 0 ? 1 : (d == 1 ? N : N * N);
                const int pt = d == 0 ? kk * N + jj : (d == 1 ? kk * N + ii : jj * N + ii);
                const double* sfd = col + (d * 6) * RS - idx * step;
                const double* slp = col + 18 * RS - idx * step;
                double dv[7] = {0, 0, 0, 0, 0, 0, 0};
                if (d == 0 && N % 2 == 0) {
                    // a line along x1 is N consecutive doubles, 16-byte aligned: 128-bit loads (same sums, same order)
                    const double2* s2 = reinterpret_cast<const double2*>(sfd);
                    const double2* l2 = reinterpret_cast<const double2*>(slp);
#pragma unroll
                    for (int m = 0; m < N; m += 2) {
                        const double d0 = s_op[C::O_DIFF + idx * N + m], d1 = s_op[C::O_DIFF + idx * N + m + 1];
#pragma unroll
                        for (int v = 0; v < 6; ++v) {
                            const double2 t = s2[(v * RS + m) / 2];
                            dv[v] += d0 * t.x;
                            dv[v] += d1 * t.y;
                        }
                        const double2 t = l2[m / 2];
                        dv[6] += s_op[C::O_DLP + idx * N + m] * t.x;
                        dv[6] += s_op[C::O_DLP + idx * N + m + 1] * t.y;
                    }
                } else {
#pragma unroll
                    for (int m = 0; m < N; ++m) {
                        const double dm = s_op[C::O_DIFF + idx * N + m];
#pragma unroll
                        for (int v = 0; v < 6; ++v) dv[v] += dm * sfd[v * RS + m * step];
                        dv[6] += s_op[C::O_DLP + idx * N + m] * slp[m * step];
                        if (d == 2) grav += s_op[C::O_HF + idx * N + m] * slp[RS + m * step];
                    }
                }
                const double cw = s_op[C::O_CORR + idx * 2], ce = s_op[C::O_CORR + idx * 2 + 1];
                const double* flo = d == 0 ? &sm.ffx[sa * (TI + 1) + sb][0][pt] : (d == 1 ? &sm.ffy[sa * TI + sb][0][pt] : &sm.ffd[le][0][pt]);
                const double* fhi = d == 0 ? &sm.ffx[sa * (TI + 1) + sb + 1][0][pt] : (d == 1 ? &sm.ffy[(sa + 1) * TI + sb][0][pt] : &sm.ffu[le][0][pt]);
#pragma unroll
                for (int v = 0; v < 5; ++v) dv[v] += cw * flo[v * F] + ce * fhi[v * F];
                // this element is the max-side (R) element of its min-side face and the L element of its max-side face
                dv[5] += cw * flo[(d == 2 ? 5 : 6) * F] + ce * fhi[5 * F];
                // d(log p) + lift of the own-side face log p = gamma (D + c_w (x) ext_w + c_e (x) ext_e) lg + const(idx)
                dv[6] = WXB_GAMMA * dv[6] + s_op[C::O_KLP + idx];
#pragma unroll
                for (int v = 0; v < 4; ++v) acc[v] += dv[v];
                const double wpres = col[(d * 6 + 5) * RS];
                wacc += dv[4] + p * dv[5] + p * wpres * dv[6];
            }
            const double inv_sg = wxbfm::fast_rcp(sg);
            double r[NVAR];
            r[0] = -inv_sg * acc[0];
            r[1] = -inv_sg * acc[1] - fo[0];
            r[2] = -inv_sg * acc[2] - fo[1];
            r[3] = -inv_sg * wacc - (fo[2] + gfac * WXB_GRAVITY * inv_sg * grav);
            r[4] = -inv_sg * acc[3];
#pragma unroll
            for (int v = 0; v < NVAR; ++v) {
                double x = r[v];
                if (JVP) {
                    x = a.dt * (x - r0[v]) * a.inv_eps;  // matvec.py:66
                    if (a.mode == WXB_JVP_RAT) x = vv[v] - 0.5 * x;
                }
                a.out[qo + v * vstride] = x;
            }
        }
        if (WXB_SW_ALIAS) __syncthreads();  // the next step stages rows where the flux fields were
    }
}

// ------------------------------------------------------------------------------------------------------
// Trace records for OTHER ranks (multi-rank handles): the same staged rows and the same form_trace as the sweep, for
// the (column, face) pairs of the send list only.  One block per (record, level) for the levels [k0, k1) of the call;
// sendbuf[rec][k - k0][field][F] (k0 = 0, k1 = nz: the layout the ghost records have on the receiving rank; the host
// pipeline sends level chunks, each chunk's records behind one another, and unpacks on arrival: wxb_unpack_ghost).
template <int N, bool JVP>
__global__ void __launch_bounds__(N* N* N) euler_edge_traces(const EvalArgs a) {
    constexpr int P = N * N * N, F = N * N;
    __shared__ double srow[NSTAGE][P];
    __shared__ double s_ext[2 * N];
    __shared__ double s_tab[WXB_FASTTAB_DOUBLES];
    const int kc = a.k1 - a.k0;
    const int rec = blockIdx.x / kc, kl = blockIdx.x - rec * kc, k = a.k0 + kl;
    const int node = threadIdx.x;
    for (int t = node; t < WXB_FASTTAB_DOUBLES; t += P) s_tab[t] = a.fasttab[t];
    if (node < 2 * N) s_ext[node] = a.ops[node];
    __syncthreads();
    const int cf = a.send[rec];
    const int c = cf >> 2, f = cf & 3;
    int s = 0;
    for (int u = 1; u < a.st.nslabs; ++u)
        if (c >= a.st.cbase[u]) s = u;
    const int cl = c - a.st.cbase[s];
    const int e = a.st.ebase[s] + k * a.st.ny[s] * a.nx + cl;
    const ElemPos ep = decode_elem(a.st, e, a.nz, a.nx, P, NVAR);
    const long long o = ep.qoff + node;
    double x[NVAR];
#pragma unroll
    for (int v = 0; v < NVAR; ++v) x[v] = a.q[o + v * ep.vstride];
    if (JVP) {
        double dv[NVAR];
#pragma unroll
        for (int v = 0; v < NVAR; ++v) dv[v] = a.v[o + v * ep.vstride];
#pragma unroll
        for (int v = 0; v < NVAR; ++v) x[v] += a.eps * dv[v];
    }
    const int sn = swz_node<N>(node);
    srow[0][sn] = x[1];
    srow[1][sn] = x[2];
    srow[2][sn] = x[3];
    srow[R_LRHO][sn] = wxbfm::fast_log(x[0], s_tab);
    srow[R_LRT][sn] = wxbfm::fast_log(x[4], s_tab);
    __syncthreads();
    if (node < F) {
        double o6[NTR];
        form_trace<N>(&srow[0][0], P, s_ext + (f & 1) * N, f >> 1, node, s_tab + WXB_EXPTAB_OFF, o6);
        double* dst = a.sendbuf + (((long long)rec * kc + kl) * NTR) * F + node;
#pragma unroll
        for (int v = 0; v < NTR; ++v) dst[v * F] = o6[v];
    }
}

// ghost[(g nz + k) rec + x] = stage[(g kc + k - k0) rec + x] for the levels of one chunk (rec = NTR F doubles)
__global__ void __launch_bounds__(256) unpack_ghost(const double* __restrict__ stage, double* __restrict__ ghost, long long total, int rec,
                                                   int kc, int k0, int nz) {
    const long long t = blockIdx.x * 256LL + threadIdx.x;
    if (t >= total) return;
    const long long r = t / rec;
    const int x = (int)(t - r * rec);
    const long long g = r / kc;
    const int kl = (int)(r - g * kc);
    ghost[(g * nz + k0 + kl) * rec + x] = stage[t];
}

template <int N>
int launch_sweep(const EvalArgs& a, cudaStream_t s) {
    using C = SCfg<N>;
    if (a.ntiles == 0 || a.k1 <= a.k0) return 0;
    constexpr int smem = (int)sizeof(SweepSmem<N>);
    if (a.mode >= 0)
        euler_sweep<N, true><<<a.ntiles, C::THREADS, smem, s>>>(a);
    else
        euler_sweep<N, false><<<a.ntiles, C::THREADS, smem, s>>>(a);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

template <int N>
int launch_edge(const EvalArgs& a, cudaStream_t s) {
    if (a.nsend == 0 || a.k1 <= a.k0) return 0;
    if (a.v)
        euler_edge_traces<N, true><<<a.nsend * (a.k1 - a.k0), N * N * N, 0, s>>>(a);
    else
        euler_edge_traces<N, false><<<a.nsend * (a.k1 - a.k0), N * N * N, 0, s>>>(a);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

template <int N>
int configure_sweep() {
    constexpr int smem = (int)sizeof(SweepSmem<N>);
    if (cudaFuncSetAttribute(euler_sweep<N, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1;
    if (cudaFuncSetAttribute(euler_sweep<N, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return -1;
    return 0;
}

}  // namespace

int wxb_sweep_configure(int n) {
    switch (n) {
        case 2: return configure_sweep<2>();
        case 3: return configure_sweep<3>();
        case 4: return configure_sweep<4>();
        case 5: return configure_sweep<5>();
        case 6: return configure_sweep<6>();
    }
    return -1;
}

void wxb_sweep_tile_shape(int n, int* tj, int* ti) {
    switch (n) {
        case 2: *tj = SCfg<2>::TJ; *ti = SCfg<2>::TI; break;
        case 3: *tj = SCfg<3>::TJ; *ti = SCfg<3>::TI; break;
        case 4: *tj = SCfg<4>::TJ; *ti = SCfg<4>::TI; break;
        case 5: *tj = SCfg<5>::TJ; *ti = SCfg<5>::TI; break;
        default: *tj = SCfg<6>::TJ; *ti = SCfg<6>::TI; break;
    }
}

int wxb_euler_sweep(const EvalArgs& a, int n, cudaStream_t s) {
    switch (n) {
        case 2: return launch_sweep<2>(a, s);
        case 3: return launch_sweep<3>(a, s);
        case 4: return launch_sweep<4>(a, s);
        case 5: return launch_sweep<5>(a, s);
        case 6: return launch_sweep<6>(a, s);
    }
    return -1;
}

int wxb_unpack_ghost(const double* stage, double* ghost, int nghost, int rec, int k0, int k1, int nz, cudaStream_t s) {
    const long long total = (long long)nghost * (k1 - k0) * rec;
    if (total <= 0) return 0;
    unpack_ghost<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(stage, ghost, total, rec, k1 - k0, k0, nz);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

int wxb_euler_edge_traces(const EvalArgs& a, int n, cudaStream_t s) {
    switch (n) {
        case 2: return launch_edge<2>(a, s);
        case 3: return launch_edge<3>(a, s);
        case 4: return launch_edge<4>(a, s);
        case 5: return launch_edge<5>(a, s);
        case 6: return launch_edge<6>(a, s);
    }
    return -1;
}
