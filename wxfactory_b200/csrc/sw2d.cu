// Shallow-water DFR right-hand side on the cubed sphere -- sm_100a kernels (FP64).
// Restructures RhsShallowWater.__compute_rhs__ (wx_factory/rhs/rhs_sw.py:58-240) into two kernels:
//   sw_traces : (h + hsurf, hu1, hu2) extrapolated to the 4 element edges            (:81-90)
//   sw_main   : pointwise fluxes (:120-131), sum-factorised interior derivative (:134-135), table-driven pull
//               of the neighbour trace with panel-edge rotation/flip (:103-117,138-150; process_topology.py:128-175),
//               topography removal on the traces (:153-155), AUSM interface flux (:170-207), lift (:210-211),
//               Coriolis / metric / topography forcing (:223-235), assembly (:238); FD-JVP epilogue (matvec.py:57-88).
// Point orders: node = j*n + i ; x1-edge pt = j ; x2-edge pt = i ; faces 0 W 1 E 2 S 3 N.
#include "wxb_internal.cuh"

namespace {

constexpr int NVAR = 3;

template <int N>
struct Cfg2 {
    static constexpr int P = N * N;
    static constexpr int F = N;
    static constexpr int EB = 256 / P;
    static constexpr int THREADS = EB * P;
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
    const int e = blockIdx.x * EB + le;
    const bool active = e < a.E;
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

// AUSM-type interface flux at one edge point; out3 = common flux of (h, hu1, hu2)
template <int N>
__device__ __forceinline__ void sw_face_point(const EvalArgs& a, const ElemPos& ep, int e, int f, int pt, double* out3,
                                              int out_stride) {
    constexpr int F = N;
    const int d = f >> 1, side = f & 1;
    const long long fstride = (long long)a.E * F;
    double own[NVAR], nb[NVAR];
    {
        const double* t = a.trace + ((long long)f * a.E + e) * F + pt;
#pragma unroll
        for (int v = 0; v < NVAR; ++v) own[v] = t[v * 4 * fstride];
    }
    {
        const int4 t = a.nbr[(long long)f * a.C + ep.c];
        const int pt2 = (t.w & 1) ? (N - 1 - pt) : pt;
        const double* src = a.trace + (long long)t.x * F + pt2;
        const long long vs = (long long)t.z * F;
#pragma unroll
        for (int v = 0; v < NVAR; ++v) nb[v] = src[v * vs];
        const int xid = t.w >> 8;
        if (xid) {
            const double* m = a.xform + ((long long)(xid - 1) * N + pt) * 4;
            const double b1 = m[0] * nb[1] + m[1] * nb[2];
            const double b2 = m[2] * nb[1] + m[3] * nb[2];
            nb[1] = b1;
            nb[2] = b2;
        }
    }
    double sg, hc1, hc2;
    long long fidx;  // index of this physical face point in the slab's face array (one component)
    long long cs;
    if (d == 0) {
        cs = (long long)ep.ny * (a.nx + 1) * F;
        fidx = ((long long)ep.jl * (a.nx + 1) + ep.i + side) * F + pt;
        const double* m = a.fmi + a.st.fmi[ep.s] + fidx;
        sg = m[0]; hc1 = m[cs]; hc2 = m[2 * cs];
    } else {
        cs = (long long)(ep.ny + 1) * a.nx * F;
        fidx = ((long long)(ep.jl + side) * a.nx + ep.i) * F + pt;
        const double* m = a.fmj + a.st.fmj[ep.s] + fidx;
        sg = m[0]; hc1 = m[cs]; hc2 = m[2 * cs];
    }
    double L[NVAR], R[NVAR];
#pragma unroll
    for (int v = 0; v < NVAR; ++v) {
        L[v] = side ? own[v] : nb[v];
        R[v] = side ? nb[v] : own[v];
    }
    if (a.has_topo) {  // rhs_sw.py:153-155; topo_itf: i-faces then j-faces, per slab [2][faces]: L-side, R-side value
        const double* ti;
        if (d == 0) ti = a.topo_itf + 2 * (a.st.fmi[ep.s] / 3) + fidx;
        else ti = a.topo_itf + 2 * (a.st.fmi[a.st.nslabs - 1] / 3 + (long long)a.st.ny[a.st.nslabs - 1] * (a.nx + 1) * F) +
                  2 * (a.st.fmj[ep.s] / 3) + fidx;
        L[0] -= ti[0];
        R[0] -= ti[cs];
    }
    const double hdd = d == 0 ? hc1 : hc2;
    const double aL = sqrt(WXB_GRAVITY * L[0] * hdd), aR = sqrt(WXB_GRAVITY * R[0] * hdd);
    const double denL = L[0] * aL, denR = R[0] * aR;
    const double mL = denL != 0.0 ? L[1 + d] / denL : 0.0;
    const double mR = denR != 0.0 ? R[1 + d] / denR : 0.0;
    const double bigM = 0.25 * ((mL + 1.0) * (mL + 1.0) - (mR - 1.0) * (mR - 1.0));
    const double cL = fmax(0.0, bigM) * aL, cR = fmin(0.0, bigM) * aR;
    const double hg = 0.5 * WXB_GRAVITY;
    const double p1L = sg * hg * hc1 * (L[0] * L[0]), p1R = sg * hg * hc1 * (R[0] * R[0]);
    const double p2L = sg * hg * hc2 * (L[0] * L[0]), p2R = sg * hg * hc2 * (R[0] * R[0]);
    out3[0] = sg * (cL * L[0] + cR * R[0]);
    out3[out_stride] = sg * (cL * L[1] + cR * R[1]) + 0.5 * ((1.0 + mL) * p1L + (1.0 - mR) * p1R);
    out3[2 * out_stride] = sg * (cL * L[2] + cR * R[2]) + 0.5 * ((1.0 + mL) * p2L + (1.0 - mR) * p2R);
}

template <int N>
__global__ void __launch_bounds__(Cfg2<N>::THREADS) sw_main(const EvalArgs a) {
    using C = Cfg2<N>;
    constexpr int P = C::P, F = C::F, EB = C::EB;
    __shared__ double s_op[C::NOPS];
    __shared__ double s_f[EB][2][NVAR][P];
    __shared__ double s_ff[EB][4][NVAR][F];
    const int le = threadIdx.x / P, node = threadIdx.x % P;
    const int slot = blockIdx.x * EB + le;
    const int e = a.elist ? (slot < a.nelist ? a.elist[slot] : -1) : (slot < a.E ? slot : -1);
    const bool active = e >= 0;
    for (int t = threadIdx.x; t < C::NOPS; t += C::THREADS) s_op[t] = a.ops[t];
    const int ii = node % N, jj = node / N;
    ElemPos ep;
    double q[NVAR], sg = 1.0, H11 = 0, H12 = 0, H22 = 0, u1 = 0, u2 = 0;
    const long long EP = (long long)a.E * P;
    long long mo = 0;
    if (active) {
        ep = decode_elem(a.st, e, 1, a.nx, P, NVAR);
        mo = (long long)e * P + node;
#pragma unroll
        for (int v = 0; v < NVAR; ++v) {
            const long long o = ep.qoff + v * ep.vstride + node;
            q[v] = a.q[o];
            if (a.v) q[v] += a.eps * a.v[o];
        }
        sg = a.sqrtg[mo];
        H11 = a.h[mo];
        H12 = a.h[EP + mo];
        H22 = a.h[2 * EP + mo];
        u1 = q[1] / q[0];
        u2 = q[2] / q[0];
        const double hsq = q[0] * q[0], hg = 0.5 * WXB_GRAVITY;
        s_f[le][0][0][node] = sg * q[1];
        s_f[le][0][1][node] = sg * (q[1] * u1 + hg * H11 * hsq);
        s_f[le][0][2][node] = sg * (q[2] * u1 + hg * H12 * hsq);
        s_f[le][1][0][node] = sg * q[2];
        s_f[le][1][1][node] = sg * (q[1] * u2 + hg * H12 * hsq);
        s_f[le][1][2][node] = sg * (q[2] * u2 + hg * H22 * hsq);
        for (int fp = node; fp < 4 * F; fp += P) {
            const int f = fp / F, pt = fp % F;
            sw_face_point<N>(a, ep, e, f, pt, &s_ff[le][f][0][pt], F);
        }
    }
    __syncthreads();
    if (!active) return;
    double acc[NVAR] = {0, 0, 0};
#pragma unroll
    for (int d = 0; d < 2; ++d) {
        const int idx = d == 0 ? ii : jj, step = d == 0 ? 1 : N, base = node - idx * step;
        const int pt = d == 0 ? jj : ii;
#pragma unroll
        for (int m = 0; m < N; ++m) {
            const double dm = s_op[C::O_DIFF + idx * N + m];
#pragma unroll
            for (int v = 0; v < NVAR; ++v) acc[v] += dm * s_f[le][d][v][base + m * step];
        }
        const double cw = s_op[C::O_CORR + idx * 2], ce = s_op[C::O_CORR + idx * 2 + 1];
#pragma unroll
        for (int v = 0; v < NVAR; ++v) acc[v] += cw * s_ff[le][2 * d][v][pt] + ce * s_ff[le][2 * d + 1][v][pt];
    }
    double dz1 = 0.0, dz2 = 0.0;
    if (a.has_topo) {
        dz1 = a.topo[EP + mo];
        dz2 = a.topo[2 * EP + mo];
    }
    const double* c = a.chris + mo;  // 1_01 1_02 1_11 1_12 2_01 2_02 2_12 2_22
    const double fo1 = 2.0 * (c[0] * q[1] + c[EP] * q[2]) + c[2 * EP] * q[1] * u1 + 2.0 * c[3 * EP] * q[1] * u2 +
                       WXB_GRAVITY * q[0] * (H11 * dz1 + H12 * dz2);
    const double fo2 = 2.0 * (c[4 * EP] * q[1] + c[5 * EP] * q[2]) + 2.0 * c[6 * EP] * q[1] * u2 + c[7 * EP] * q[2] * u2 +
                       WXB_GRAVITY * q[0] * (H12 * dz1 + H22 * dz2);
    const double inv_sg = 1.0 / sg;
    double r[NVAR];
    r[0] = -inv_sg * acc[0];
    r[1] = -inv_sg * acc[1] - fo1;
    r[2] = -inv_sg * acc[2] - fo2;
#pragma unroll
    for (int v = 0; v < NVAR; ++v) {
        const long long o = ep.qoff + v * ep.vstride + node;
        double x = r[v];
        if (a.mode >= 0) {
            x = a.dt * (x - a.rhs0[o]) / a.eps;
            if (a.mode == WXB_JVP_RAT) x = a.v[o] - 0.5 * x;
        }
        a.out[o] = x;
    }
}

template <int N>
int launch_traces2(const EvalArgs& a, cudaStream_t s) {
    const int blocks = (a.E + Cfg2<N>::EB - 1) / Cfg2<N>::EB;
    sw_traces<N><<<blocks, Cfg2<N>::THREADS, 0, s>>>(a);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}
template <int N>
int launch_main2(const EvalArgs& a, cudaStream_t s) {
    const int ne = a.elist ? a.nelist : a.E;
    if (ne == 0) return 0;
    const int blocks = (ne + Cfg2<N>::EB - 1) / Cfg2<N>::EB;
    sw_main<N><<<blocks, Cfg2<N>::THREADS, 0, s>>>(a);
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
