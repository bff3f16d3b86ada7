// Shallow-water DFR right-hand side on the cubed sphere -- sm_100a kernels (FP64).
// Restructures RhsShallowWater.__compute_rhs__ (wx_factory/rhs/rhs_sw.py:58-240) into two kernels:
//   sw_traces : (h + hsurf, hu1, hu2) extrapolated to the 4 element edges            (:81-90)
//   sw_main   : pointwise fluxes (:120-131), sum-factorised interior derivative (:134-135), table-driven pull
//               of the neighbour trace with panel-edge rotation/flip (:103-117,138-150; process_topology.py:128-175),
//               topography removal on the traces (:153-155), AUSM interface flux (:170-207), lift (:210-211),
//               Coriolis / metric / topography forcing (:223-235), assembly (:238); FD-JVP epilogue (matvec.py:57-88).
// Point orders: node = j*n + i ; x1-edge pt = j ; x2-edge pt = i ; faces 0 W 1 E 2 S 3 N.
#include "wxb_internal.cuh"

#ifndef WXB_SW_MINB
#define WXB_SW_MINB 4
#endif
#ifndef WXB_SW_THREADS
#define WXB_SW_THREADS 256
#endif

namespace {

constexpr int NVAR = 3;

template <int N>
struct Cfg2 {
    static constexpr int P = N * N;
    static constexpr int F = N;
    static constexpr int EB = WXB_SW_THREADS / P;
    static constexpr int THREADS = EB * P;
    static constexpr int MINB = WXB_SW_MINB;  // resident blocks per SM sw_main is compiled for
    static constexpr int O_DIFF = 2 * N, O_CORR = 2 * N + N * N;
    static constexpr int NOPS = 4 * N + N * N;
};

template <int N>
__global__ void __launch_bounds__(Cfg2<N>::THREADS) sw_traces(const EvalArgs a) {
    using C = Cfg2<N>;
    constexpr int P = C::P, F = C::F, EB = C::EB;
    __shared__ double s_ext[2 * N];
    __shared__ double s_q[EB][NVAR][P];
    const int le = threadIdx.x / P, node = threadIdx.x % P;
    const int slot = blockIdx.x * EB + le;
    const int e = a.elist ? (slot < a.nelist ? a.elist[slot] : -1) : (slot < a.E ? slot : -1);
    const bool active = e >= 0;
    if (threadIdx.x < 2 * N) s_ext[threadIdx.x] = a.ops[threadIdx.x];
    if (active) {
        const ElemPos ep = decode_elem(a.st, e, 1, a.nx, P, NVAR);
#pragma unroll
        for (int v = 0; v < NVAR; ++v) {
            const long long o = ep.qoff + v * ep.vstride + node;
            double x = a.q[o];
            if (a.v) x += a.eps * a.v[o];
            if (v == 0 && a.has_topo) x += a.topo[(long long)e * P + node];
            s_q[le][v][node] = x;
        }
    }
    __syncthreads();
    if (!active) return;
    const long long fstride = (long long)a.E * F;
    for (int ln = node; ln < 2 * F; ln += P) {
        const int d = ln / F, pt = ln % F;
        const int base = d == 0 ? pt * N : pt, step = d == 0 ? 1 : N;
        double lo[NVAR] = {0, 0, 0}, hi[NVAR] = {0, 0, 0};
#pragma unroll
        for (int m = 0; m < N; ++m) {
#pragma unroll
            for (int v = 0; v < NVAR; ++v) {
                const double x = s_q[le][v][base + m * step];
                lo[v] += s_ext[m] * x;
                hi[v] += s_ext[N + m] * x;
            }
        }
        double* t = a.trace + (long long)e * F + pt;
#pragma unroll
        for (int v = 0; v < NVAR; ++v) {
            t[(v * 4 + 2 * d) * fstride] = lo[v];
            t[(v * 4 + 2 * d + 1) * fstride] = hi[v];
        }
    }
}

// ------------------------------------------------------------------------------------------------------
// Per-element bookkeeping, computed once per element by one thread and kept in shared memory.
struct SwInfo {
    long long qoff, vstride;  // (var 0, this element, node 0) in the state; distance between variables
    const double* fm[2];      // (comp 0, min-side edge, pt 0) of this element in fm_i / fm_j
    long long fm_cs[2];       // component stride of the slab's block
    long long fm_side[2];     // min-side -> max-side edge
    const double* ti[2];      // two-sided topography on the min-side edge (L value; R value fm_cs further), or null
    int e, c;
};

template <int N>
__device__ __forceinline__ void sw_make_info(const EvalArgs& a, int e, SwInfo& o) {
    constexpr int F = N, P = N * N;
    const ElemPos ep = decode_elem(a.st, e, 1, a.nx, P, NVAR);
    o.qoff = ep.qoff;
    o.vstride = ep.vstride;
    o.e = e;
    o.c = ep.c;
    const long long nx = a.nx, ny = ep.ny;
    const long long fi = ((long long)ep.jl * (nx + 1) + ep.i) * F;
    const long long fj = ((long long)ep.jl * nx + ep.i) * F;
    o.fm[0] = a.fmi + a.st.fmi[ep.s] + fi;
    o.fm_cs[0] = ny * (nx + 1) * F;
    o.fm_side[0] = F;
    o.fm[1] = a.fmj + a.st.fmj[ep.s] + fj;
    o.fm_cs[1] = (ny + 1) * nx * F;
    o.fm_side[1] = nx * F;
    o.ti[0] = o.ti[1] = nullptr;
    if (a.has_topo) {  // topo_itf: i-faces then j-faces, per slab [2][faces]: L-side, R-side value (rhs_sw.py:153-155)
        const int last = a.st.nslabs - 1;
        o.ti[0] = a.topo_itf + 2 * (a.st.fmi[ep.s] / 3) + fi;
        o.ti[1] = a.topo_itf + 2 * (a.st.fmi[last] / 3 + (long long)a.st.ny[last] * (nx + 1) * F) + 2 * (a.st.fmj[ep.s] / 3) + fj;
    }
}

// everything one edge point reads from global memory; requested together, used later
struct SwFaceIn {
    double own[NVAR], nb[NVAR], sg, hc1, hc2, tL, tR;
};

template <int N>
__device__ __forceinline__ void sw_face_load(const EvalArgs& a, const SwInfo& ei, const int4 t, int f, int pt, SwFaceIn& r) {
    constexpr int F = N;
    const int d = f >> 1, side = f & 1;
    const long long fstride = (long long)a.E * F;
    const double* to = a.trace + ((long long)f * a.E + ei.e) * F + pt;
    const int pt2 = (t.w & 1) ? (N - 1 - pt) : pt;
    const double* src = a.trace + (long long)t.x * F + pt2;
    const long long vs = (long long)t.z * F;
    const long long so = side ? ei.fm_side[d] : 0;
    const double* m = ei.fm[d] + so + pt;
    const long long cs = ei.fm_cs[d];
#pragma unroll
    for (int v = 0; v < NVAR; ++v) r.own[v] = to[v * 4 * fstride];
#pragma unroll
    for (int v = 0; v < NVAR; ++v) r.nb[v] = src[v * vs];
    r.sg = m[0];
    r.hc1 = m[cs];
    r.hc2 = m[2 * cs];
    r.tL = r.tR = 0.0;
    if (a.has_topo) {
        const double* ti = ei.ti[d] + so + pt;
        r.tL = ti[0];
        r.tR = ti[cs];
    }
}

// AUSM-type interface flux at one edge point (rhs_sw.py:138-207); out3 = common flux of (h, hu1, hu2)
template <int N>
__device__ __forceinline__ void sw_face_flux(const EvalArgs& a, const int4 t, int f, int pt, SwFaceIn& r, double* out3, int out_stride) {
    const int d = f >> 1, side = f & 1;
    const int xid = t.w >> 8;
    if (xid) {  // panel edge: rotate the facing side's momenta into this panel's frame (process_topology.py:128-175)
        const double* m = a.xform + ((long long)(xid - 1) * N + pt) * 4;
        const double b1 = m[0] * r.nb[1] + m[1] * r.nb[2];
        const double b2 = m[2] * r.nb[1] + m[3] * r.nb[2];
        r.nb[1] = b1;
        r.nb[2] = b2;
    }
    double L[NVAR], R[NVAR];
#pragma unroll
    for (int v = 0; v < NVAR; ++v) {
        L[v] = side ? r.own[v] : r.nb[v];
        R[v] = side ? r.nb[v] : r.own[v];
    }
    if (a.has_topo) {
        L[0] -= r.tL;
        R[0] -= r.tR;
    }
    const double sg = r.sg, hc1 = r.hc1, hc2 = r.hc2;
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
    out3[0] = sg * (cL * L[0] + cR * R[0]);
    out3[out_stride] = sg * (cL * L[1] + cR * R[1]) + 0.5 * ((1.0 + mL) * p1L + (1.0 - mR) * p1R);
    out3[2 * out_stride] = sg * (cL * L[2] + cR * R[2]) + 0.5 * ((1.0 + mL) * p2L + (1.0 - mR) * p2R);
}

// main kernel: EB elements per block, one thread per solution point.
//   setup    one thread per element decodes it, four look its facing rows up
//   edges    the block's 4*EB*F edge points are spread over ALL its threads, edge-major, so that the own traces,
//            the facing traces of interior elements and the edge metric are read as runs of EB*F doubles
//   nodes    every operand of the node (state, metric, Christoffel, base RHS of the JVP) requested up front
//   finish   sum-factorised derivative + lift, assembly, FD-JVP epilogue
template <int N, bool JVP>
__global__ void __launch_bounds__(Cfg2<N>::THREADS, Cfg2<N>::MINB) sw_main(const EvalArgs a) {
    using C = Cfg2<N>;
    constexpr int P = C::P, F = C::F, EB = C::EB, THREADS = C::THREADS;
    constexpr int NFP = 4 * EB * F, ROUNDS = (NFP + THREADS - 1) / THREADS;
    __shared__ double s_op[C::NOPS];
    __shared__ double s_f[2 * NVAR][EB * P];
    __shared__ double s_ff[EB][4 * NVAR][F];
    __shared__ SwInfo s_ei[EB];
    __shared__ int4 s_nr[EB][4];
    const int le = threadIdx.x / P, node = threadIdx.x % P;
    const int slot = blockIdx.x * EB + le;
    const int e = a.elist ? (slot < a.nelist ? a.elist[slot] : -1) : (slot < a.E ? slot : -1);
    const bool active = e >= 0;
    for (int t = threadIdx.x; t < C::NOPS; t += THREADS) s_op[t] = a.ops[t];
    if (node == 0) {
        if (active) sw_make_info<N>(a, e, s_ei[le]);
        else s_ei[le].e = -1;
    }
    if (active && node >= P - 4) {  // one level: local column number == local element number
        const int f = node - (P - 4);
        s_nr[le][f] = a.nbr[(long long)f * a.C + e];
    }
    __syncthreads();

    // ---- edges ----
    {
        SwFaceIn fin[ROUNDS];
        int fe[ROUNDS], ff[ROUNDS], fpt[ROUNDS];
#pragma unroll
        for (int r = 0; r < ROUNDS; ++r) {
            const int t = threadIdx.x + r * THREADS;
            fe[r] = -1;
            if (t < NFP) {
                const int f = t / (EB * F), rem = t % (EB * F);
                const int el = rem / F;
                ff[r] = f;
                fpt[r] = rem % F;
                if (s_ei[el].e >= 0) {
                    fe[r] = el;
                    sw_face_load<N>(a, s_ei[el], s_nr[el][f], f, fpt[r], fin[r]);
                }
            }
        }
#pragma unroll
        for (int r = 0; r < ROUNDS; ++r)
            if (fe[r] >= 0) sw_face_flux<N>(a, s_nr[fe[r]][ff[r]], ff[r], fpt[r], fin[r], &s_ff[fe[r]][ff[r] * NVAR][fpt[r]], F);
    }

    // ---- nodes ----
    const int ii = node % N, jj = node / N;
    const long long EP = (long long)a.E * P;
    double q[NVAR], r0[NVAR], vv[NVAR], sg = 1.0, H11 = 0, H12 = 0, H22 = 0, u1 = 0, u2 = 0, c[8], dz1 = 0.0, dz2 = 0.0;
    long long qo = 0, vst = 0;
    if (active) {
        const SwInfo& ei = s_ei[le];
        qo = ei.qoff + node;
        vst = ei.vstride;
        const long long mo = (long long)e * P + node;
#pragma unroll
        for (int v = 0; v < NVAR; ++v) q[v] = a.q[qo + v * vst];
        if (JVP) {
#pragma unroll
            for (int v = 0; v < NVAR; ++v) vv[v] = a.v[qo + v * vst];
        }
        sg = a.sqrtg[mo];
        H11 = a.h[mo];
        H12 = a.h[EP + mo];
        H22 = a.h[2 * EP + mo];
#pragma unroll
        for (int m = 0; m < 8; ++m) c[m] = a.chris[m * EP + mo];  // 1_01 1_02 1_11 1_12 2_01 2_02 2_12 2_22
        if (a.has_topo) {
            dz1 = a.topo[EP + mo];
            dz2 = a.topo[2 * EP + mo];
        }
        if (JVP) {
#pragma unroll
            for (int v = 0; v < NVAR; ++v) r0[v] = a.rhs0[qo + v * vst];
#pragma unroll
            for (int v = 0; v < NVAR; ++v) q[v] += a.eps * vv[v];
        }
        const double ih = 1.0 / q[0];  // one reciprocal serves both velocities
        u1 = q[1] * ih;
        u2 = q[2] * ih;
        const double hsq = q[0] * q[0], hg = 0.5 * WXB_GRAVITY;
        const int col = le * P + node;
        s_f[0][col] = sg * q[1];
        s_f[1][col] = sg * (q[1] * u1 + hg * H11 * hsq);
        s_f[2][col] = sg * (q[2] * u1 + hg * H12 * hsq);
        s_f[3][col] = sg * q[2];
        s_f[4][col] = sg * (q[1] * u2 + hg * H12 * hsq);
        s_f[5][col] = sg * (q[2] * u2 + hg * H22 * hsq);
    }
    __syncthreads();
    if (!active) return;
    double acc[NVAR] = {0, 0, 0};
#pragma unroll
    for (int d = 0; d < 2; ++d) {
        const int idx = d == 0 ? ii : jj, step = d == 0 ? 1 : N, base = le * P + node - idx * step;
        const int pt = d == 0 ? jj : ii;
#pragma unroll
        for (int m = 0; m < N; ++m) {
            const double dm = s_op[C::O_DIFF + idx * N + m];
#pragma unroll
            for (int v = 0; v < NVAR; ++v) acc[v] += dm * s_f[d * NVAR + v][base + m * step];
        }
        const double cw = s_op[C::O_CORR + idx * 2], ce = s_op[C::O_CORR + idx * 2 + 1];
#pragma unroll
        for (int v = 0; v < NVAR; ++v) acc[v] += cw * s_ff[le][2 * d * NVAR + v][pt] + ce * s_ff[le][(2 * d + 1) * NVAR + v][pt];
    }
    const double fo1 = 2.0 * (c[0] * q[1] + c[1] * q[2]) + c[2] * q[1] * u1 + 2.0 * c[3] * q[1] * u2 +
                       WXB_GRAVITY * q[0] * (H11 * dz1 + H12 * dz2);
    const double fo2 = 2.0 * (c[4] * q[1] + c[5] * q[2]) + 2.0 * c[6] * q[1] * u2 + c[7] * q[2] * u2 +
                       WXB_GRAVITY * q[0] * (H12 * dz1 + H22 * dz2);
    const double inv_sg = 1.0 / sg;
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
        a.out[qo + v * vst] = x;
    }
}

template <int N>
int launch_traces2(const EvalArgs& a, cudaStream_t s) {
    const int ne = a.elist ? a.nelist : a.E;
    if (ne == 0) return 0;
    const int blocks = (ne + Cfg2<N>::EB - 1) / Cfg2<N>::EB;
    sw_traces<N><<<blocks, Cfg2<N>::THREADS, 0, s>>>(a);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
template <int N>
int launch_main2(const EvalArgs& a, cudaStream_t s) {
    const int ne = a.elist ? a.nelist : a.E;
    if (ne == 0) return 0;
    const int blocks = (ne + Cfg2<N>::EB - 1) / Cfg2<N>::EB;
    if (a.mode >= 0)
        sw_main<N, true><<<blocks, Cfg2<N>::THREADS, 0, s>>>(a);
    else
        sw_main<N, false><<<blocks, Cfg2<N>::THREADS, 0, s>>>(a);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

}  // namespace

int wxb_sw_traces(const EvalArgs& a, int n, cudaStream_t s) {
    switch (n) {
        case 2: return launch_traces2<2>(a, s);
        case 3: return launch_traces2<3>(a, s);
        case 4: return launch_traces2<4>(a, s);
        case 5: return launch_traces2<5>(a, s);
        case 6: return launch_traces2<6>(a, s);
    }
    return -1;
}
int wxb_sw_main(const EvalArgs& a, int n, cudaStream_t s) {
    switch (n) {
        case 2: return launch_main2<2>(a, s);
        case 3: return launch_main2<3>(a, s);
        case 4: return launch_main2<4>(a, s);
        case 5: return launch_main2<5>(a, s);
        case 6: return launch_main2<6>(a, s);
    }
    return -1;
}
