// Shallow-water DFR right-hand side on the cubed sphere -- the fused tile kernel (sm_100a, FP64).
//
// One launch per evaluation and NO trace array (the 2-D counterpart of euler_sweep.cu).  A block owns a TJ x TI tile of
// elements, one thread per solution point.  wx_factory/rhs/rhs_sw.py:58-240 in four phases:
//
//   stage    the tile's own elements and the 2(TJ+TI) elements around it are read from the state (q, or q + eps*v for
//            the FD Jacobian-vector product); (h + hsurf, hu1, hu2) go to shared memory                       (:81-90)
//   edges    every PHYSICAL edge of the tile is evaluated once, one thread per edge point: both sides' traces are
//            extrapolated from the staged rows (a halo element through the pull table, with the panel-edge reversal /
//            basis change of wx_factory/process_topology.py:128-175; on a multi-rank handle a ghost record), the
//            topography is removed (:153-155) and the AUSM flux formed (:170-207)
//   nodes    pointwise fluxes (:120-131) -> shared memory
//   finish   sum-factorised derivative (:134-135) + lift (:210-211), Coriolis / metric / topography forcing (:223-235),
//            assembly (:238), FD-JVP epilogue (wx_factory/solvers/matvec.py:57-88)
//
// Against the two-kernel path of sw2d.cu: no trace array is written or read back (96 B per node at n = 3 on top of the
// 160 algorithmic ones), the state is read once from DRAM (halo re-reads hit L2), interior edges are computed once
// instead of twice, and all element bookkeeping (decode, neighbour look-up) is done once per TILE: with 9 nodes per
// element that bookkeeping was 40 % of sw_main's instructions.
//
// The arithmetic of every value is the same, in the same order, as in sw2d.cu (the extrapolation sums, the flux, the
// derivative): the two paths agree bit for bit.
#include "fastmath.cuh"
#include "wxb_internal.cuh"

#ifndef WXB_SWT_TJ
#define WXB_SWT_TJ 4
#endif
#ifndef WXB_SWT_TI
#define WXB_SWT_TI 8
#endif
#ifndef WXB_SWT_MINB
#define WXB_SWT_MINB 3
#endif

namespace {

constexpr int NVAR = 3;

template <int N>
struct TCfg {
    static constexpr int P = N * N, F = N;
    // tile: ~256-320 threads
    static constexpr int TJ = N <= 3 ? WXB_SWT_TJ : (N == 4 ? 4 : 2);
    static constexpr int TI = N <= 3 ? WXB_SWT_TI : (N == 4 ? 4 : (N == 5 ? 4 : 4));
    static constexpr int NE = TJ * TI, NH = 2 * (TJ + TI);
    static constexpr int THREADS = NE * P;
    static constexpr int NXF = TJ * (TI + 1), NYF = (TJ + 1) * TI, NFACE = NXF + NYF;
    static constexpr int O_DIFF = 2 * N, O_CORR = 2 * N + N * N;
    static constexpr int NOPS = 4 * N + N * N;
};

struct THalo {
    long long off0;     // kind 1: (var 0, node 0) of the neighbour element in the state; kind 2: field 0, point 0 of the ghost record
    long long vstride;  // kind 1: distance between variables in the neighbour's slab; kind 2: distance between fields
    long long eoff;     // kind 1: neighbour's local element * P (for the topography added to h)
    int flags;          // bit0 = reversed along the edge, bits 8.. = 1 + xform row
    short fq, kind;     // neighbour's face that touches us; 0 none, 1 staged from the state, 2 ghost record
};

struct TFace {
    const double* fm;   // edge metric, component 0, point 0
    long long cs;       // component stride
    const double* ti;   // two-sided topography on the edge (L value; R value cs further), or null
    short lo, hi;       // min-side / max-side element: own slot, 100 + halo slot, -1 = edge does not exist
};

template <int N>
struct TileSmem {
    using C = TCfg<N>;
    double so[NVAR][C::NE * C::P];   // own elements: h + hsurf, hu1, hu2
    double sh[NVAR][C::NH * C::P];   // halo elements
    double fl[2 * NVAR][C::NE * C::P];
    double ff[C::NFACE][NVAR][C::F]; // common flux per physical edge
    double op[C::NOPS];
    THalo hd[C::NH];
};

// trace of one edge point from an element's staged rows (rows `rs` apart): the same sums as sw_traces (sw2d.cu)
template <int N>
__device__ __forceinline__ void tile_trace(const double* __restrict__ mp, int rs, const double* __restrict__ ext, int d, int pt, double out[NVAR]) {
    const int base = d == 0 ? pt * N : pt, step = d == 0 ? 1 : N;
#pragma unroll
    for (int v = 0; v < NVAR; ++v) out[v] = 0.0;
#pragma unroll
    for (int m = 0; m < N; ++m) {
#pragma unroll
        for (int v = 0; v < NVAR; ++v) out[v] += ext[m] * mp[v * rs + base + m * step];
    }
}

// AUSM-type flux at one physical edge point from its min-side (L) and max-side (R) traces (rhs_sw.py:170-207);
// same expressions as sw_face_flux (sw2d.cu)
__device__ __forceinline__ void ausm_lr(int d, const double L[NVAR], const double R[NVAR], double sg, double hc1, double hc2, double out[NVAR]) {
    const double hdd = d == 0 ? hc1 : hc2;
    const double nL = d == 0 ? L[1] : L[2], nR = d == 0 ? R[1] : R[2];
    const double aL = sqrt(WXB_GRAVITY * L[0] * hdd), aR = sqrt(WXB_GRAVITY * R[0] * hdd);
    const double denL = L[0] * aL, denR = R[0] * aR;
    const double mL = denL != 0.0 ? nL / denL : 0.0;
    const double mR = denR != 0.0 ? nR / denR : 0.0;
    const double bigM = 0.25 * ((mL + 1.0) * (mL + 1.0) - (mR - 1.0) * (mR - 1.0));
    const double cL = fmax(0.0, bigM) * aL, cR = fmin(0.0, bigM) * aR;
    const double hg = 0.5 * WXB_GRAVITY;
    const double p1L = sg * hg * hc1 * (L[0] * L[0]), p1R = sg * hg * hc1 * (R[0] * R[0]);
    const double p2L = sg * hg * hc2 * (L[0] * L[0]), p2R = sg * hg * hc2 * (R[0] * R[0]);
    out[0] = sg * (cL * L[0] + cR * R[0]);
    out[1] = sg * (cL * L[1] + cR * R[1]) + 0.5 * ((1.0 + mL) * p1L + (1.0 - mR) * p1R);
    out[2] = sg * (cL * L[2] + cR * R[2]) + 0.5 * ((1.0 + mL) * p2L + (1.0 - mR) * p2R);
}

template <int N, bool JVP>
__global__ void __launch_bounds__(TCfg<N>::THREADS, WXB_SWT_MINB) sw_tile(const EvalArgs a) {
    using C = TCfg<N>;
    constexpr int P = C::P, F = C::F, TJ = C::TJ, TI = C::TI, NE = C::NE, NH = C::NH, NT = C::THREADS;
    constexpr int RSO = NE * P, RSH = NH * P;
    __shared__ TileSmem<N> sm;
    const int tid = threadIdx.x, le = tid / P, node = tid - le * P;

    for (int t = tid; t < C::NOPS; t += NT) sm.op[t] = a.ops[t];
    const int tile = a.tlist ? a.tlist[blockIdx.x] : (int)blockIdx.x;
    const int nx = a.nx;
    const int4 ti4 = a.tinfo[tile];  // slab, first row, first column (built on the host at create)
    const int s = ti4.x, tj0 = ti4.y, ti0 = ti4.z;
    const int ny = a.st.ny[s];
    const int aj = min(TJ, ny - tj0), ai = min(TI, nx - ti0);
    const long long vstride = (long long)ny * nx * P;
    const int sa = le / TI, sb = le - sa * TI;
    const bool active = sa < aj && sb < ai;
    const long long el = (long long)(tj0 + sa) * nx + ti0 + sb;  // element within the slab
    const long long e = a.st.ebase[s] + el;                       // ... and in the handle's element order
    const long long qo = (long long)NVAR * P * a.st.ebase[s] + el * P + node;
    const long long EP = (long long)a.E * P;

    // halo slot h -> which edge of the tile, which own element it touches, and the neighbour's position in the slab
    auto halo_geom = [&](int h, int& face, int& jn, int& in) -> bool {
        int ha, hb;
        bool ok;
        if (h < TJ) { face = 0; ha = h; hb = 0; ok = ha < aj; }
        else if (h < 2 * TJ) { face = 1; ha = h - TJ; hb = ai - 1; ok = ha < aj; }
        else if (h < 2 * TJ + TI) { face = 2; ha = 0; hb = h - 2 * TJ; ok = hb < ai; }
        else { face = 3; ha = aj - 1; hb = h - 2 * TJ - TI; ok = hb < ai; }
        jn = tj0 + ha + (face == 2 ? -1 : (face == 3 ? 1 : 0));
        in = ti0 + hb + (face == 0 ? -1 : (face == 1 ? 1 : 0));
        return ok;
    };
    // A neighbour inside the slab is the adjacent element of the same panel: identity transform, address known without the
    // table (wx_factory/process_topology.py only acts on panel edges).  Only tiles on a slab edge take the table path.
    if (tid < NH) {
        int face, jn, in;
        const bool ok = halo_geom(tid, face, jn, in);
        THalo d;
        d.off0 = d.vstride = d.eoff = 0;
        d.flags = 0;
        d.fq = d.kind = 0;
        if (ok && jn >= 0 && jn < ny && in >= 0 && in < nx) {
            const long long en = (long long)jn * nx + in;
            d.off0 = (long long)NVAR * P * a.st.ebase[s] + en * P;
            d.vstride = vstride;
            d.eoff = (a.st.ebase[s] + en) * P;
            d.fq = (short)(face ^ 1);
            d.kind = 3;  // staged from the state, loads already issued by the halo threads below
        } else if (ok) {
            const int c = a.st.cbase[s] + (jn - (face == 2 ? -1 : (face == 3 ? 1 : 0))) * nx + in - (face == 0 ? -1 : (face == 1 ? 1 : 0));
            const int4 t = a.nbr[(long long)face * a.C + c];  // base, level stride, field stride, flags (rows of F doubles)
            d.flags = t.w;
            if ((long long)t.x < (long long)NVAR * 4 * a.E) {  // a local element: [field][face][E] numbering
                const int en = t.x % a.E;
                d.fq = (short)(t.x / a.E);
                const ElemPos ep = decode_elem(a.st, en, 1, nx, P, NVAR);
                d.off0 = ep.qoff;
                d.vstride = ep.vstride;
                d.eoff = (long long)en * P;
                d.kind = 1;
            } else {  // a ghost record [rec][field][1][F] received from another rank
                d.off0 = (long long)t.x * F;
                d.vstride = (long long)t.z * F;
                d.kind = 2;
            }
        }
        sm.hd[tid] = d;
    }

    // ---- every global load the common case needs is requested here, before the first barrier: the block is short-lived
    // (one tile), so each dependent round trip it exposes costs ~1 us of a few-us life ----
    // (1) this thread's own node
    double q[NVAR] = {1.0, 0.0, 0.0}, vv[NVAR] = {0.0, 0.0, 0.0};
    double r0[NVAR] = {0.0, 0.0, 0.0}, sg = 1.0, H11 = 0, H12 = 0, H22 = 0, c[8], dz1 = 0.0, dz2 = 0.0, ht = 0.0;
    if (active) {
        const long long mo = e * P + node;
#pragma unroll
        for (int v = 0; v < NVAR; ++v) q[v] = a.q[qo + v * vstride];
        if (JVP) {
#pragma unroll
            for (int v = 0; v < NVAR; ++v) vv[v] = a.v[qo + v * vstride];
        }
        if (a.has_topo) {
            ht = a.topo[mo];
            dz1 = a.topo[EP + mo];
            dz2 = a.topo[2 * EP + mo];
        }
        sg = a.sqrtg[mo];
        H11 = a.h[mo];
        H12 = a.h[EP + mo];
        H22 = a.h[2 * EP + mo];
#pragma unroll
        for (int m = 0; m < 8; ++m) c[m] = a.chris[m * EP + mo];  // 1_01 1_02 1_11 1_12 2_01 2_02 2_12 2_22
        if (JVP) {
#pragma unroll
            for (int v = 0; v < NVAR; ++v) r0[v] = a.rhs0[qo + v * vstride];
        }
    }
    // (2) the halo nodes whose element lies inside the slab
    constexpr int HR = (NH * P + NT - 1) / NT;
    double xh[HR][NVAR], vh[HR][JVP ? NVAR : 1], th[HR];
    bool hin[HR];
#pragma unroll
    for (int r = 0; r < HR; ++r) {
        const int t = tid + r * NT;
        hin[r] = false;
        th[r] = 0.0;
        if (t < NH * P) {
            const int h = t / P, hn = t - h * P;
            int face, jn, in;
            if (halo_geom(h, face, jn, in) && jn >= 0 && jn < ny && in >= 0 && in < nx) {
                hin[r] = true;
                const long long en = (long long)jn * nx + in;
                const long long o = (long long)NVAR * P * a.st.ebase[s] + en * P + hn;
#pragma unroll
                for (int v = 0; v < NVAR; ++v) xh[r][v] = a.q[o + v * vstride];
                if (JVP) {
#pragma unroll
                    for (int v = 0; v < NVAR; ++v) vh[r][v] = a.v[o + v * vstride];
                }
                if (a.has_topo) th[r] = a.topo[(a.st.ebase[s] + en) * P + hn];
            }
        }
    }
    // (3) the edge metric of this thread's edge points
    constexpr int FR = (C::NFACE * F + NT - 1) / NT;
    double em[FR][5];
    short elo[FR], ehi[FR];
#pragma unroll
    for (int r = 0; r < FR; ++r) {
        const int fp = tid + r * NT;
        elo[r] = ehi[r] = -1;
        if (fp < C::NFACE * F) {
            const int f = fp / F, pt = fp - f * F;
            const double* fm = nullptr;
            const double* ti = nullptr;
            long long cs = 0;
            if (f < C::NXF) {
                const int fa = f / (TI + 1), pos = f - fa * (TI + 1);
                if (fa < aj && pos <= ai) {
                    elo[r] = pos > 0 ? fa * TI + pos - 1 : 100 + fa;
                    ehi[r] = pos < ai ? fa * TI + pos : 100 + TJ + fa;
                    const long long fi = ((long long)(tj0 + fa) * (nx + 1) + ti0 + pos) * F;
                    fm = a.fmi + a.st.fmi[s] + fi;
                    cs = (long long)ny * (nx + 1) * F;
                    if (a.has_topo) ti = a.topo_itf + 2 * (a.st.fmi[s] / 3) + fi;
                }
            } else {
                const int g = f - C::NXF, pos = g / TI, fb = g - pos * TI;
                if (fb < ai && pos <= aj) {
                    elo[r] = pos > 0 ? (pos - 1) * TI + fb : 100 + 2 * TJ + fb;
                    ehi[r] = pos < aj ? pos * TI + fb : 100 + 2 * TJ + TI + fb;
                    const long long fj = ((long long)(tj0 + pos) * nx + ti0 + fb) * F;
                    fm = a.fmj + a.st.fmj[s] + fj;
                    cs = (long long)(ny + 1) * nx * F;
                    if (a.has_topo) {
                        const int last = a.st.nslabs - 1;
                        ti = a.topo_itf + 2 * (a.st.fmi[last] / 3 + (long long)a.st.ny[last] * (nx + 1) * F) + 2 * (a.st.fmj[s] / 3) + fj;
                    }
                }
            }
            if (elo[r] >= 0) {
                em[r][0] = fm[pt];
                em[r][1] = fm[cs + pt];
                em[r][2] = fm[2 * cs + pt];
                em[r][3] = em[r][4] = 0.0;
                if (a.has_topo) {
                    em[r][3] = ti[pt];
                    em[r][4] = ti[cs + pt];
                }
            }
        }
    }

    // ---- stage, and everything of the node that needs no neighbour: flux fields -> shared memory, forcing -> 2 registers
    // (the 20 values loaded above die here, long before the edge phase) ----
    double fo1 = 0.0, fo2 = 0.0, inv_sg = 1.0;
    if (active) {
        if (JVP) {
#pragma unroll
            for (int v = 0; v < NVAR; ++v) q[v] += a.eps * vv[v];
        }
        double h0 = q[0];
        if (a.has_topo) h0 += ht;
        sm.so[0][tid] = h0;
        sm.so[1][tid] = q[1];
        sm.so[2][tid] = q[2];
        const double ih = 1.0 / q[0];  // one reciprocal serves both velocities
        const double u1 = q[1] * ih, u2 = q[2] * ih;
        const double hsq = q[0] * q[0], hg = 0.5 * WXB_GRAVITY;
        sm.fl[0][tid] = sg * q[1];
        sm.fl[1][tid] = sg * (q[1] * u1 + hg * H11 * hsq);
        sm.fl[2][tid] = sg * (q[2] * u1 + hg * H12 * hsq);
        sm.fl[3][tid] = sg * q[2];
        sm.fl[4][tid] = sg * (q[1] * u2 + hg * H12 * hsq);
        sm.fl[5][tid] = sg * (q[2] * u2 + hg * H22 * hsq);
        fo1 = 2.0 * (c[0] * q[1] + c[1] * q[2]) + c[2] * q[1] * u1 + 2.0 * c[3] * q[1] * u2 + WXB_GRAVITY * q[0] * (H11 * dz1 + H12 * dz2);
        fo2 = 2.0 * (c[4] * q[1] + c[5] * q[2]) + 2.0 * c[6] * q[1] * u2 + c[7] * q[2] * u2 + WXB_GRAVITY * q[0] * (H12 * dz1 + H22 * dz2);
        inv_sg = 1.0 / sg;
    }
#pragma unroll
    for (int r = 0; r < HR; ++r) {
        if (hin[r]) {
            const int t = tid + r * NT;
#pragma unroll
            for (int v = 0; v < NVAR; ++v) sm.sh[v][t] = JVP ? xh[r][v] + a.eps * vh[r][v] : xh[r][v];
            if (a.has_topo) sm.sh[0][t] += th[r];
        }
    }
    __syncthreads();  // descriptors, own rows, in-slab halo rows
    // tiles on a slab edge: the neighbours the table points at (another slab of this rank)
    for (int t = tid; t < NH * P; t += NT) {
        const int h = t / P, hn = t - h * P;
        const THalo& hd = sm.hd[h];
        if (hd.kind != 1) continue;
        const long long o = hd.off0 + hn;
        double x[NVAR];
#pragma unroll
        for (int v = 0; v < NVAR; ++v) x[v] = a.q[o + v * hd.vstride];
        if (JVP) {
#pragma unroll
            for (int v = 0; v < NVAR; ++v) x[v] += a.eps * a.v[o + v * hd.vstride];
        }
        if (a.has_topo) x[0] += a.topo[hd.eoff + hn];
#pragma unroll
        for (int v = 0; v < NVAR; ++v) sm.sh[v][t] = x[v];
    }
    __syncthreads();

    // ---- edges: one thread per physical edge point ----
#pragma unroll
    for (int r = 0; r < FR; ++r) {
        const int fp = tid + r * NT;
        if (elo[r] < 0) continue;
        const int f = fp / F, pt = fp - f * F;
        const int d = f < C::NXF ? 0 : 1;
        double L[NVAR], R[NVAR];
        auto side = [&](int code, int own_side, double out[NVAR]) {
            if (code < 100) {
                tile_trace<N>(&sm.so[0][code * P], RSO, sm.op + own_side * N, d, pt, out);
                return;
            }
            const THalo& hd = sm.hd[code - 100];
            const int pt2 = (hd.flags & 1) ? (N - 1 - pt) : pt;
            if (hd.kind == 2) {
                const double* g = a.trace + hd.off0 + pt2;
#pragma unroll
                for (int v = 0; v < NVAR; ++v) out[v] = g[v * hd.vstride];
            } else {
                tile_trace<N>(&sm.sh[0][(code - 100) * P], RSH, sm.op + (hd.fq & 1) * N, hd.fq >> 1, pt2, out);
            }
            const int xid = hd.flags >> 8;
            if (xid) {  // the neighbour's (hu1, hu2) in this panel's basis (process_topology.py:138-175)
                const double* x = a.xform + ((long long)(xid - 1) * N + pt) * 4;
                const double b1 = x[0] * out[1] + x[1] * out[2];
                const double b2 = x[2] * out[1] + x[3] * out[2];
                out[1] = b1;
                out[2] = b2;
            }
        };
        side(elo[r], 1, L);
        side(ehi[r], 0, R);
        if (a.has_topo) {
            L[0] -= em[r][3];
            R[0] -= em[r][4];
        }
        double cf[NVAR];
        ausm_lr(d, L, R, em[r][0], em[r][1], em[r][2], cf);
#pragma unroll
        for (int v = 0; v < NVAR; ++v) sm.ff[f][v][pt] = cf[v];
    }

    __syncthreads();
    if (!active) return;

    // ---- finish ----
    const int ii = node % N, jj = node / N;
    double acc[NVAR] = {0, 0, 0};
#pragma unroll
    for (int d = 0; d < 2; ++d) {
        const int idx = d == 0 ? ii : jj, step = d == 0 ? 1 : N, base = tid - idx * step;
        const int pt = d == 0 ? jj : ii;
#pragma unroll
        for (int m = 0; m < N; ++m) {
            const double dm = sm.op[C::O_DIFF + idx * N + m];
#pragma unroll
            for (int v = 0; v < NVAR; ++v) acc[v] += dm * sm.fl[d * NVAR + v][base + m * step];
        }
        const double cw = sm.op[C::O_CORR + idx * 2], ce = sm.op[C::O_CORR + idx * 2 + 1];
        const int flo = d == 0 ? sa * (TI + 1) + sb : C::NXF + sa * TI + sb;
        const int fhi = d == 0 ? flo + 1 : flo + TI;
#pragma unroll
        for (int v = 0; v < NVAR; ++v) acc[v] += cw * sm.ff[flo][v][pt] + ce * sm.ff[fhi][v][pt];
    }
    double r[NVAR];
    r[0] = -inv_sg * acc[0];
    r[1] = -inv_sg * acc[1] - fo1;
    r[2] = -inv_sg * acc[2] - fo2;
#pragma unroll
    for (int v = 0; v < NVAR; ++v) {
        double x = r[v];
        if (JVP) {
            x = a.dt * (x - r0[v]) * a.inv_eps;  // matvec.py:66 divides by eps: same to one rounding
            if (a.mode == WXB_JVP_RAT) x = vv[v] - 0.5 * x;
        }
        a.out[qo + v * vstride] = x;
    }
}

// Trace records for OTHER ranks (multi-rank handles): one block per record; sendbuf[rec][field][1][F], the same sums as
// the tile kernel's own side.
template <int N, bool JVP>
__global__ void __launch_bounds__(N* N) sw_edge_traces(const EvalArgs a) {
    constexpr int P = N * N, F = N;
    __shared__ double srow[NVAR][P];
    __shared__ double s_ext[2 * N];
    const int rec = blockIdx.x, node = threadIdx.x;
    if (node < 2 * N) s_ext[node] = a.ops[node];
    const int cf = a.send[rec];
    const int e = cf >> 2, f = cf & 3;  // one level: local column number == local element number
    const ElemPos ep = decode_elem(a.st, e, 1, a.nx, P, NVAR);
#pragma unroll
    for (int v = 0; v < NVAR; ++v) {
        const long long o = ep.qoff + v * ep.vstride + node;
        double x = a.q[o];
        if (JVP) x += a.eps * a.v[o];
        if (v == 0 && a.has_topo) x += a.topo[(long long)e * P + node];
        srow[v][node] = x;
    }
    __syncthreads();
    if (node < F) {
        double o3[NVAR];
        tile_trace<N>(&srow[0][0], P, s_ext + (f & 1) * N, f >> 1, node, o3);
#pragma unroll
        for (int v = 0; v < NVAR; ++v) a.sendbuf[((long long)rec * NVAR + v) * F + node] = o3[v];
    }
}

template <int N>
int launch_tile(const EvalArgs& a, cudaStream_t s) {
    if (a.ntiles == 0) return 0;
    if (a.mode >= 0)
        sw_tile<N, true><<<a.ntiles, TCfg<N>::THREADS, 0, s>>>(a);
    else
        sw_tile<N, false><<<a.ntiles, TCfg<N>::THREADS, 0, s>>>(a);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

template <int N>
int launch_tile_edge(const EvalArgs& a, cudaStream_t s) {
    if (a.nsend == 0) return 0;
    if (a.v)
        sw_edge_traces<N, true><<<a.nsend, N * N, 0, s>>>(a);
    else
        sw_edge_traces<N, false><<<a.nsend, N * N, 0, s>>>(a);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // namespace

void wxb_sw_tile_shape(int n, int* tj, int* ti) {
    switch (n) {
        case 2: *tj = TCfg<2>::TJ; *ti = TCfg<2>::TI; break;
        case 3: *tj = TCfg<3>::TJ; *ti = TCfg<3>::TI; break;
        case 4: *tj = TCfg<4>::TJ; *ti = TCfg<4>::TI; break;
        case 5: *tj = TCfg<5>::TJ; *ti = TCfg<5>::TI; break;
        default: *tj = TCfg<6>::TJ; *ti = TCfg<6>::TI; break;
    }
}

int wxb_sw_tile(const EvalArgs& a, int n, cudaStream_t s) {
    switch (n) {
        case 2: return launch_tile<2>(a, s);
        case 3: return launch_tile<3>(a, s);
        case 4: return launch_tile<4>(a, s);
        case 5: return launch_tile<5>(a, s);
        case 6: return launch_tile<6>(a, s);
    }
    return -1;
}

int wxb_sw_tile_edge_traces(const EvalArgs& a, int n, cudaStream_t s) {
    switch (n) {
        case 2: return launch_tile_edge<2>(a, s);
        case 3: return launch_tile_edge<3>(a, s);
        case 4: return launch_tile_edge<4>(a, s);
        case 5: return launch_tile_edge<5>(a, s);
        case 6: return launch_tile_edge<6>(a, s);
    }
    return -1;
}
