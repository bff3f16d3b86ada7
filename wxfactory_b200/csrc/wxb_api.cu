// C ABI of libwxb200.so (see include/wxb200.h for the contract and the reference interfaces replaced).
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <dlfcn.h>
#include <new>
#include <vector>

#include "fastmath.cuh"
#include "wxb_internal.cuh"

namespace {

// orders whose Euler handles run the fused column-sweep kernel by default (the others run the traces + main pair)
#ifndef WXB_SWEEP_DEFAULT
#define WXB_SWEEP_DEFAULT(n) ((n) == 2 || (n) == 4)
#endif

char g_create_err[512] = "";
void comm_free(wxb_handle* h);

void set_err(wxb_handle* h, const char* fmt, ...) {
    char* dst = h ? h->err : g_create_err;
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(dst, 512, fmt, ap);
    va_end(ap);
}

#define CU(call)                                                                              \
    do {                                                                                      \
        cudaError_t _e = (call);                                                              \
        if (_e != cudaSuccess) {                                                              \
            set_err(h, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return _e == cudaErrorMemoryAllocation ? WXB_ENOMEM : WXB_ECUDA;                  \
        }                                                                                     \
    } while (0)

template <typename T>
int upload(wxb_handle* h, T** dst, const T* src, size_t count) {
    *dst = nullptr;
    if (count == 0) return WXB_OK;
    if (!src) {
        set_err(h, "missing static array (%zu elements expected)", count);
        return WXB_EINVAL;
    }
    CU(cudaMalloc((void**)dst, count * sizeof(T)));
    CU(cudaMemcpy(*dst, src, count * sizeof(T), cudaMemcpyHostToDevice));
    return WXB_OK;
}

#define UP(dst, src, count)                          \
    do {                                             \
        int _r = upload(h, &(dst), (src), (count));  \
        if (_r != WXB_OK) return _r;                 \
    } while (0)

int build(wxb_handle* h, const wxb_config* cfg, const wxb_static_arrays* A) {
    const bool euler = cfg->equations == WXB_EQ_EULER3D;
    const int n = cfg->n;
    h->n = n;
    h->nvar = euler ? 5 : 3;
    h->ntr = euler ? 6 : 3;  // trace fields per face point: Euler adds log(rho*theta) (see euler3d.cu)
    h->P = euler ? n * n * n : n * n;
    h->F = euler ? n * n : n;
    h->nfaces = euler ? 6 : 4;
    const int nz = euler ? cfg->nz : 1;
    const int nx = cfg->nx;
    const int ncomp = euler ? 4 : 3;
    SlabTab& st = h->st;
    memset(&st, 0, sizeof(st));
    st.nslabs = cfg->nslabs;
    long long oi = 0, oj = 0, ok = 0;
    int eb = 0, cb = 0;
    for (int s = 0; s < cfg->nslabs; ++s) {
        const int ny = cfg->slab_ny[s];
        if (ny <= 0) {
            set_err(h, "slab %d has ny=%d", s, ny);
            return WXB_EINVAL;
        }
        st.ny[s] = ny;
        st.ebase[s] = eb;
        st.cbase[s] = cb;
        st.fmi[s] = oi;
        st.fmj[s] = oj;
        st.fmk[s] = ok;
        eb += nz * ny * nx;
        cb += ny * nx;
        oi += (long long)ncomp * nz * ny * (nx + 1) * h->F;
        oj += (long long)ncomp * nz * (ny + 1) * nx * h->F;
        ok += (long long)ncomp * (nz + 1) * ny * nx * h->F;
    }
    for (int s = cfg->nslabs; s <= WXB_MAX_SLABS; ++s) {
        st.ebase[s] = eb;
        st.cbase[s] = cb;
    }
    h->E = eb;
    h->C = cb;
    const size_t EP = (size_t)h->E * h->P;

    const int nops = euler ? 5 * n + 3 * n * n : 4 * n + n * n;
    std::vector<double> ops(nops, 0.0);
    if (!A->ext_w || !A->ext_e || !A->diff || !A->corr || (euler && !A->hfilt)) {
        set_err(h, "missing operator block");
        return WXB_EINVAL;
    }
    memcpy(&ops[0], A->ext_w, n * sizeof(double));
    memcpy(&ops[n], A->ext_e, n * sizeof(double));
    memcpy(&ops[2 * n], A->diff, n * n * sizeof(double));
    memcpy(&ops[2 * n + n * n], A->corr, 2 * n * sizeof(double));
    if (euler) {
        memcpy(&ops[4 * n + n * n], A->hfilt, n * n * sizeof(double));
        // The reference differentiates log p = log p0 + gamma*lg (lg = log(Rd/p0 * rho theta)) at the nodes and lifts
        // the own-side face value log p0 + gamma*(sum_m ext[m] log(rho theta)_m + log(Rd/p0)) (rhs_dfr.py:106-139,
        // pde_euler_cubesphere.py:124,158-160).  Both are linear in the line's lg values, so the kernel applies ONE
        // matrix  M = D + c_w (x) ext_w + c_e (x) ext_e  to lg and adds the constant that the float64 operator
        // entries leave over (rows of D + corrections do not sum to exactly 0, extrapolation rows not to exactly 1):
        //   K[i] = log p0 * (sum_m D[i][m] + c_w[i] + c_e[i]) + gamma*log(Rd/p0) * (c_w[i](1 - sum ext_w) + c_e[i](1 - sum ext_e))
        // formed here in extended precision, i.e. the exactly-rounded value of what the reference's formulas define.
        const long double lp0 = logl(100000.0L), lrd = logl(287.05L / 100000.0L), gam = (long double)(double)WXB_GAMMA;
        long double sw = 0.0L, se = 0.0L;
        for (int m = 0; m < n; ++m) {
            sw += (long double)A->ext_w[m];
            se += (long double)A->ext_e[m];
        }
        double* M = &ops[4 * n + 2 * n * n];
        double* K = M + n * n;
        for (int i = 0; i < n; ++i) {
            const long double cw = A->corr[2 * i], ce = A->corr[2 * i + 1];
            long double rs = cw + ce;
            for (int m = 0; m < n; ++m) {
                rs += (long double)A->diff[i * n + m];
                M[i * n + m] = (double)((long double)A->diff[i * n + m] + cw * (long double)A->ext_w[m] + ce * (long double)A->ext_e[m]);
            }
            K[i] = (double)(lp0 * rs + gam * lrd * (cw * (1.0L - sw) + ce * (1.0L - se)));
        }
    }
    UP(h->d_ops, ops.data(), (size_t)nops);

    if (euler) {
        // one block, element-major: [E][35][P] with rows Christoffel 0..26 | sqrtG | h 0..5 | 1/z_eta.  A thread
        // addresses all 35 values of its node with one base pointer and compile-time offsets, and an element's
        // 35 rows (17.9 KB at n = 4) are one contiguous piece of HBM (one bulk copy in the TMA-staged variant).
        if (!A->sqrtg || !A->h || !A->chris || !A->inv_dzdeta) {
            set_err(h, "missing interior metric array");
            return WXB_EINVAL;
        }
        CU(cudaMalloc((void**)&h->d_metric, 35 * EP * sizeof(double)));
        h->d_chris = h->d_metric;
        h->d_sqrtg = h->d_h = h->d_idz = nullptr;  // rows of the block, not separate arrays
        {
            double* tmp = nullptr;  // one host row [E*P] at a time, interleaved on the device
            CU(cudaMalloc((void**)&tmp, EP * sizeof(double)));
            const size_t rowb = (size_t)h->P * sizeof(double);
            for (int r = 0; r < 35; ++r) {
                const double* src = r < 27 ? A->chris + (size_t)r * EP
                                  : r == 27 ? A->sqrtg
                                  : r < 34 ? A->h + (size_t)(r - 28) * EP
                                           : A->inv_dzdeta;
                cudaError_t e1 = cudaMemcpy(tmp, src, EP * sizeof(double), cudaMemcpyHostToDevice);
                cudaError_t e2 = e1 == cudaSuccess ? cudaMemcpy2D(h->d_metric + (size_t)r * h->P, 35 * rowb, tmp, rowb, rowb, (size_t)h->E, cudaMemcpyDeviceToDevice) : e1;
                if (e2 != cudaSuccess) {
                    cudaFree(tmp);
                    set_err(h, "metric upload failed: %s", cudaGetErrorString(e2));
                    return WXB_ECUDA;
                }
            }
            cudaFree(tmp);
        }
    } else {
        UP(h->d_sqrtg, A->sqrtg, EP);
        UP(h->d_h, A->h, 3 * EP);
        UP(h->d_chris, A->chris, 8 * EP);
    }
    if (euler && cfg->has_damping) UP(h->d_damp, A->damp, 4 * EP);
    if (!euler && cfg->has_topo) {
        UP(h->d_topo, A->topo, 3 * EP);
        // hsurf on faces, two values (own side, other side) per face point: i-faces then j-faces
        UP(h->d_topo_itf, A->topo_itf, (size_t)(oi / ncomp + oj / ncomp) * 2);
    }
    UP(h->d_fmi, A->fm_i, (size_t)oi);
    UP(h->d_fmj, A->fm_j, (size_t)oj);
    if (euler) UP(h->d_fmk, A->fm_k, (size_t)ok);
    if (euler) {
        // interface metric: the host hands [4 components][faces][F] per slab; euler_main reads the 4 components of
        // a face point together, so each slab block is transposed once to [faces][4][F] (512 contiguous bytes per
        // face at n = 4, component offsets known at compile time)
        double* dirs[3] = {h->d_fmi, h->d_fmj, h->d_fmk};
        const long long tot[3] = {oi, oj, ok};
        for (int d = 0; d < 3; ++d) {
            if (tot[d] == 0) continue;
            double* tmp = nullptr;
            CU(cudaMalloc((void**)&tmp, (size_t)tot[d] * sizeof(double)));
            cudaError_t e1 = cudaMemcpy(tmp, dirs[d], (size_t)tot[d] * sizeof(double), cudaMemcpyDeviceToDevice);
            for (int s = 0; s < cfg->nslabs && e1 == cudaSuccess; ++s) {
                const long long ny = st.ny[s];
                const long long nfaces = d == 0 ? (long long)nz * ny * (nx + 1) : d == 1 ? (long long)nz * (ny + 1) * nx : (long long)(nz + 1) * ny * nx;
                const long long off = d == 0 ? st.fmi[s] : d == 1 ? st.fmj[s] : st.fmk[s];
                const size_t rowb = (size_t)h->F * sizeof(double);
                for (int c = 0; c < 4 && e1 == cudaSuccess; ++c)
                    e1 = cudaMemcpy2D(dirs[d] + off + (size_t)c * h->F, 4 * rowb, tmp + off + (size_t)c * nfaces * h->F, rowb, rowb, (size_t)nfaces, cudaMemcpyDeviceToDevice);
            }
            cudaFree(tmp);
            if (e1 != cudaSuccess) {
                set_err(h, "interface metric transpose failed: %s", cudaGetErrorString(e1));
                return WXB_ECUDA;
            }
        }
    }
    {
        int4* tmp = nullptr;
        int r = upload(h, (int32_t**)&tmp, A->nbr, (size_t)4 * h->C * 4);
        if (r != WXB_OK) return r;
        h->d_nbr = tmp;
    }
    if (euler) {  // element-major trace layout: the fields of a facing row are one row apart (partition.py "element")
        for (size_t t = 0; t < (size_t)4 * h->C; ++t)
            if (A->nbr[t * 4 + 2] != 1) {
                set_err(h, "Euler handles need the element-major trace layout (field stride 1 in the neighbour table, got %d)", A->nbr[t * 4 + 2]);
                return WXB_EINVAL;
            }
    }
    UP(h->d_xform, A->xform, (size_t)cfg->nxform * n * 4);
    UP(h->d_send, A->send, (size_t)cfg->nsend);

    // Which kernels serve the handle: the fused kernel (Euler: euler_sweep.cu, WXB_EULER_PATH=sweep, default; shallow water:
    // sw_tile.cu, WXB_SW_PATH=tile, default) or the traces + main pair (euler3d.cu / sw2d.cu, "split").  One path serves
    // BOTH operators of a handle: the FD Jacobian-vector product subtracts the cached RHS(Q) from RHS(Q + eps v), and the
    // rounding of the two evaluations only cancels (amplified by 1/eps otherwise) when both come from the same kernels --
    // a mixed mode (RHS fused, JVP pair) was measured faster and broke the Krylov iteration counts of Epi(2) on dcmip31.
    {
        const char* pe = getenv(euler ? "WXB_EULER_PATH" : "WXB_SW_PATH");
        // default per order and size: the path that measured faster (profiles/r02_sweep_orders.jsonl, r02_order_variants.log)
        bool fused_default = euler ? WXB_SWEEP_DEFAULT(n) : true;
        if (euler && fused_default) {
            // The sweep kernel gives a block a tile of columns and walks up it: a handle with few tiles cannot fill the
            // GPU for long (the reference authors' 60^3-node benchmark shape at n = 4 is 384 tiles of 15 levels for
            // 296 resident blocks: 0.242 ms against 0.157 ms on the element-parallel pair).  Below 4 full waves of
            // blocks the pair serves the handle.  (Blocks per SM as compiled in euler_sweep.cu: WXB_SW_MINB<n>.)
            int tj = 1, ti = 1;
            wxb_sweep_tile_shape(n, &tj, &ti);
            long long tiles = 0;
            for (int s = 0; s < cfg->nslabs; ++s) tiles += (long long)((st.ny[s] + tj - 1) / tj) * ((nx + ti - 1) / ti);
            const int blocks_per_sm = n == 2 ? 4 : (n == 6 ? 1 : 2);
            int sms = 148;
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, h->device);
            if (tiles < 4LL * sms * blocks_per_sm) fused_default = false;
        }
        h->split = pe ? strcmp(pe, "split") == 0 : !fused_default;
        h->sweep = !h->split;
    }
    // trace array: local part (SW [ntr][nfaces][E][F], Euler [E][nfaces][ntr][F]), then the ghost records
    // (SW [rec][ntr][1][F], Euler [rec][nz][ntr][F]); the same number of doubles either way.  The sweep path has no
    // local traces: only the ghost tail exists, addressed through the same offsets.
    h->trace_local = (int64_t)h->ntr * h->nfaces * h->E * h->F;
    const int64_t ghost = (int64_t)cfg->nghost * h->ntr * nz * h->F;
    {
        const int64_t alloc = (h->split ? h->trace_local : 0) + ghost + 1;
        CU(cudaMalloc((void**)&h->d_trace_alloc, (size_t)alloc * sizeof(double)));
        CU(cudaMemset(h->d_trace_alloc, 0, (size_t)alloc * sizeof(double)));
        h->d_trace = h->split ? h->d_trace_alloc : h->d_trace_alloc - h->trace_local;
    }
    if (cfg->nsend > 0) CU(cudaMalloc((void**)&h->d_sendbuf, (size_t)cfg->nsend * h->ntr * nz * h->F * sizeof(double)));
    if (euler) {  // log / exp tables of fastmath.cuh
        std::vector<double> tab(WXB_FASTTAB_DOUBLES);
        wxbfm::fill_tables(tab.data());
        UP(h->d_fasttab, tab.data(), tab.size());
    }
    if (h->sweep) {
        // tiles of TJ x TI columns, slab after slab, row-major; a tile is "outer" if one of its columns reads a ghost
        int tj = 1, ti = 1;
        if (euler) wxb_sweep_tile_shape(n, &tj, &ti);
        else wxb_sw_tile_shape(n, &tj, &ti);
        std::vector<int> tin, tout, tinfo;
        const int32_t ghost_base = (int32_t)(h->trace_local / h->F);
        int tile = 0;
        for (int s = 0; s < cfg->nslabs; ++s) {
            const int ny = st.ny[s];
            for (int j0 = 0; j0 < ny; j0 += tj)
                for (int i0 = 0; i0 < nx; i0 += ti, ++tile) {
                    bool g = false;
                    for (int j = j0; j < ny && j < j0 + tj; ++j)
                        for (int i = i0; i < nx && i < i0 + ti; ++i) {
                            const int c = st.cbase[s] + j * nx + i;
                            for (int f = 0; f < 4; ++f) g = g || (A->nbr[((size_t)f * h->C + c) * 4] >= ghost_base);
                        }
                    (g ? tout : tin).push_back(tile);
                    for (int x : {s, j0, i0, 0}) tinfo.push_back(x);
                }
        }
        h->nt_all = tile;
        {
            int* tmp = nullptr;
            int r = upload(h, &tmp, tinfo.data(), tinfo.size());
            if (r != WXB_OK) return r;
            h->d_tinfo = (int4*)tmp;
        }
        h->nt_inner = (int)tin.size();
        h->nt_outer = (int)tout.size();
        if (cfg->nghost > 0) {
            UP(h->d_tiles_inner, tin.data(), tin.size());
            UP(h->d_tiles_outer, tout.data(), tout.size());
        }
        if (euler && wxb_sweep_configure(n) != 0) {
            set_err(h, "cannot configure the sweep kernel's shared memory: %s", cudaGetErrorString(cudaGetLastError()));
            return WXB_ECUDA;
        }
    }

    // inner / outer element lists: an element is "outer" if one of its horizontal faces reads a ghost record
    if (cfg->nghost > 0 && h->split) {
        std::vector<int> inner, outer;
        const int32_t ghost_base = (int32_t)(h->trace_local / h->F);
        for (int s = 0; s < cfg->nslabs; ++s) {
            const int ny = st.ny[s];
            for (int k = 0; k < nz; ++k)
                for (int cl = 0; cl < ny * nx; ++cl) {
                    const int c = st.cbase[s] + cl;
                    bool g = false;
                    for (int f = 0; f < 4; ++f) g = g || (A->nbr[((size_t)f * h->C + c) * 4] >= ghost_base);
                    (g ? outer : inner).push_back(st.ebase[s] + k * ny * nx + cl);
                }
        }
        h->n_inner = (int)inner.size();
        h->n_outer = (int)outer.size();
        UP(h->d_elist_inner, inner.data(), inner.size());
        UP(h->d_elist_outer, outer.data(), outer.size());
    }
    for (int r = 0; r < WXB_TIMING_RING; ++r)
        for (int i = 0; i < 3; ++i) CU(cudaEventCreate(&h->ev_ring[r][i]));
    h->ev = h->ev_ring[0];
    if (euler && h->split && wxb_euler_configure(n) != 0) {
        set_err(h, "cannot configure the main kernel's shared memory: %s", cudaGetErrorString(cudaGetLastError()));
        return WXB_ECUDA;
    }
    return WXB_OK;
}

// the fused kernel or the traces + main pair (one path per handle, see build())
bool use_fused(const wxb_handle* h, bool /*jvp*/) { return h->sweep; }

void fill_args(wxb_handle* h, EvalArgs& a, const double* q, const double* v, const double* rhs0, double dt, double eps,
               int mode, double* out) {
    memset(&a, 0, sizeof(a));
    a.q = q;
    a.v = v;
    a.eps = v ? eps : 0.0;
    a.inv_eps = (v && eps != 0.0) ? 1.0 / eps : 0.0;
    a.rhs0 = rhs0;
    a.dt = dt;
    a.mode = mode;
    a.out = out;
    a.trace = h->d_trace;
    a.qp = h->d_qp;
    a.sendbuf = h->d_sendbuf;
    a.sqrtg = h->d_sqrtg;
    a.h = h->d_h;
    a.chris = h->d_chris;
    a.idz = h->d_idz;
    a.damp = h->d_damp;
    a.topo = h->d_topo;
    a.fmi = h->d_fmi;
    a.fmj = h->d_fmj;
    a.fmk = h->d_fmk;
    a.topo_itf = h->d_topo_itf;
    a.nbr = h->d_nbr;
    a.xform = h->d_xform;
    a.send = h->d_send;
    a.ops = h->d_ops;
    a.elist = nullptr;
    a.nelist = 0;
    a.E = h->E;
    a.C = h->C;
    a.nz = h->cfg.equations == WXB_EQ_EULER3D ? h->cfg.nz : 1;
    a.nx = h->cfg.nx;
    a.nsend = h->cfg.nsend;
    a.has_damp = h->cfg.has_damping;
    a.has_topo = h->cfg.has_topo;
    a.tau0 = h->cfg.damp_tau0;
    a.fasttab = h->d_fasttab;
    a.tinfo = h->d_tinfo;
    a.tlist = nullptr;
    a.ntiles = h->nt_all;
    a.k0 = 0;
    a.k1 = a.nz;
    a.st = h->st;
}

int check_mode(wxb_handle* h, const double* q, const double* v, const double* rhs0, double eps, int mode, double* out) {
    if (!q || !out) {
        set_err(h, "null state/output pointer");
        return WXB_EINVAL;
    }
    if (mode != -1 && mode != WXB_JVP_FD && mode != WXB_JVP_RAT) {
        set_err(h, "unknown mode %d", mode);
        return WXB_EINVAL;
    }
    if (mode >= 0 && (!v || !rhs0 || eps == 0.0)) {
        set_err(h, "JVP needs v, rhs0 and a non-zero eps");
        return WXB_EINVAL;
    }
    return WXB_OK;
}

int do_traces(wxb_handle* h, const double* q, const double* v, double eps, cudaStream_t s) {
    CU(cudaSetDevice(h->device));
    if (h->cfg.equations == WXB_EQ_EULER3D && h->cfg.advection_only) {
        h->traces_done = true;
        return WXB_OK;
    }
    const bool fused = use_fused(h, v != nullptr);
    const bool euler_eq = h->cfg.equations == WXB_EQ_EULER3D;
    if (v && !h->d_qp && !fused) {  // first JVP on this handle: scratch for the perturbed state
        CU(cudaMalloc((void**)&h->d_qp, (size_t)wxb_state_size(h) * sizeof(double)));
    }
    EvalArgs a;
    fill_args(h, a, q, v, nullptr, 0.0, eps, -1, nullptr);
    h->ev = h->ev_ring[h->nevals % WXB_TIMING_RING];
    CU(cudaEventRecord(h->ev[0], s));
    int r;
    if (fused) {  // no trace array: only the records other ranks need, straight into the send buffer
        r = euler_eq ? wxb_euler_edge_traces(a, h->n, s) : wxb_sw_tile_edge_traces(a, h->n, s);
    } else {
        r = h->cfg.equations == WXB_EQ_EULER3D ? wxb_euler_traces(a, h->n, s) : wxb_sw_traces(a, h->n, s);
        if (r == 0) r = wxb_pack(a, h->ntr, h->F, h->nfaces, h->cfg.equations == WXB_EQ_EULER3D, s);
    }
    if (r != 0) {
        set_err(h, "trace kernel launch failed: %s", cudaGetErrorString(cudaGetLastError()));
        return WXB_ECUDA;
    }
    CU(cudaEventRecord(h->ev[1], s));
    h->traces_done = true;
    return WXB_OK;
}

int do_finish(wxb_handle* h, const double* q, const double* v, const double* rhs0, double dt, double eps, int mode,
              int part, double* out, cudaStream_t s) {
    CU(cudaSetDevice(h->device));
    if (h->cfg.equations == WXB_EQ_EULER3D && h->cfg.advection_only) {
        // rhs_dfr.py:307-313: every row is zeroed -> RHS == 0 (and so is the FD difference of two zero RHS)
        if (part != 2) {
            if (mode == WXB_JVP_RAT)
                CU(cudaMemcpyAsync(out, v, (size_t)wxb_state_size(h) * sizeof(double), cudaMemcpyDeviceToDevice, s));
            else
                CU(cudaMemsetAsync(out, 0, (size_t)wxb_state_size(h) * sizeof(double), s));
        }
        return WXB_OK;
    }
    EvalArgs a;
    fill_args(h, a, q, v, rhs0, dt, eps, mode, out);
    const bool fused = use_fused(h, mode >= 0);
    if (part == 1 || part == 2) {
        if (h->cfg.nghost == 0) {
            if (part == 2) return WXB_OK;  // everything is inner
        } else if (fused) {
            a.tlist = part == 1 ? h->d_tiles_inner : h->d_tiles_outer;
            a.ntiles = part == 1 ? h->nt_inner : h->nt_outer;
        } else {
            a.elist = part == 1 ? h->d_elist_inner : h->d_elist_outer;
            a.nelist = part == 1 ? h->n_inner : h->n_outer;
        }
    } else if (part != 0) {
        set_err(h, "part must be 0, 1 or 2");
        return WXB_EINVAL;
    }
    int r = fused ? (h->cfg.equations == WXB_EQ_EULER3D ? wxb_euler_sweep(a, h->n, s) : wxb_sw_tile(a, h->n, s))
                     : (h->cfg.equations == WXB_EQ_EULER3D ? wxb_euler_main(a, h->n, s) : wxb_sw_main(a, h->n, s));
    if (r != 0) {
        set_err(h, "main kernel launch failed: %s", cudaGetErrorString(cudaGetLastError()));
        return WXB_ECUDA;
    }
    if (part != 1) {
        CU(cudaEventRecord(h->ev[2], s));
        h->ev_mode[h->nevals % WXB_TIMING_RING] = mode;
        h->nevals++;
        h->timed = true;
    }
    return WXB_OK;
}

}  // namespace

extern "C" {

int wxb_version(void) { return WXB_VERSION; }

const char* wxb_last_error(const wxb_handle* h) { return h ? h->err : g_create_err; }

int wxb_create(const wxb_config* cfg, const wxb_static_arrays* arrays, int device, wxb_handle** out) {
    wxb_handle* h = nullptr;
    if (!cfg || !arrays || !out) {
        set_err(h, "null argument");
        return WXB_EINVAL;
    }
    *out = nullptr;
    if (cfg->equations != WXB_EQ_EULER3D && cfg->equations != WXB_EQ_SHALLOW_WATER) {
        set_err(h, "unknown equation set %d", cfg->equations);
        return WXB_EINVAL;
    }
    if (cfg->n < 2 || cfg->n > WXB_MAX_N) {
        set_err(h, "num_solpts %d outside 2..%d", cfg->n, WXB_MAX_N);
        return WXB_EINVAL;
    }
    if (cfg->nslabs < 1 || cfg->nslabs > WXB_MAX_SLABS || cfg->nx < 1 ||
        (cfg->equations == WXB_EQ_EULER3D && cfg->nz < 1)) {
        set_err(h, "bad grid sizes (nslabs=%d nx=%d nz=%d)", cfg->nslabs, cfg->nx, cfg->nz);
        return WXB_EINVAL;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_err(h, "no CUDA device available: this library has no CPU fallback");
        cudaGetLastError();
        return WXB_ECUDA;
    }
    if (device < 0 || device >= ndev) {
        set_err(h, "device %d out of range (%d devices)", device, ndev);
        return WXB_EINVAL;
    }
    wxb_handle* nh = new (std::nothrow) wxb_handle();
    if (!nh) return WXB_ENOMEM;
    memset(nh, 0, sizeof(*nh));
    nh->cfg = *cfg;
    nh->device = device;
    if (cudaSetDevice(device) != cudaSuccess) {
        set_err(h, "cudaSetDevice(%d) failed", device);
        delete nh;
        return WXB_ECUDA;
    }
    int r = build(nh, cfg, arrays);
    if (r != WXB_OK) {
        snprintf(g_create_err, sizeof(g_create_err), "%s", nh->err);
        wxb_destroy(nh);
        return r;
    }
    *out = nh;
    return WXB_OK;
}

int wxb_destroy(wxb_handle* h) {
    if (!h) return WXB_OK;
    cudaSetDevice(h->device);
    if (h->d_metric) h->d_sqrtg = h->d_h = h->d_chris = h->d_idz = nullptr;  // views into d_metric
    void* ptrs[] = {h->d_ops,  h->d_sqrtg, h->d_h,        h->d_chris, h->d_idz,         h->d_damp,        h->d_topo, h->d_metric,
                    h->d_fmi,  h->d_fmj,   h->d_fmk,      h->d_topo_itf, h->d_xform,    h->d_nbr,         h->d_send,
                    h->d_elist_inner,      h->d_elist_outer, h->d_trace_alloc, h->d_sendbuf, h->d_fasttab, h->d_tiles_inner, h->d_tiles_outer, h->d_tinfo, h->d_qp,  h->d_q,           h->d_v,
                    h->d_r0,   h->d_out, h->d_lq, h->d_lr, h->d_clist, h->d_expf, h->d_recv_stage};
    for (void* p : ptrs)
        if (p) cudaFree(p);
    for (int r = 0; r < WXB_TIMING_RING; ++r)
        for (int i = 0; i < 3; ++i)
            if (h->ev_ring[r][i]) cudaEventDestroy(h->ev_ring[r][i]);
    comm_free(h);
    if (h->pipe_ready) {
        for (int c = 0; c < h->nchunk; ++c) {
            cudaEventDestroy(h->ev_in[c]);
            cudaEventDestroy(h->ev_main[c]);
            if (h->ev_packed_c[c]) cudaEventDestroy(h->ev_packed_c[c]);
            if (h->ev_halo_c[c]) cudaEventDestroy(h->ev_halo_c[c]);
        }
        cudaStreamDestroy(h->s_in);
        cudaStreamDestroy(h->s_cmp);
        cudaStreamDestroy(h->s_out);
    }
    delete h;
    return WXB_OK;
}

int wxb_handle_fused(const wxb_handle* h) { return h ? (h->sweep ? 1 : 0) : WXB_EINVAL; }

int64_t wxb_state_size(const wxb_handle* h) { return h ? (int64_t)h->nvar * h->E * h->P : 0; }
int64_t wxb_send_size(const wxb_handle* h) {
    return h ? (int64_t)h->cfg.nsend * h->ntr * (h->cfg.equations == WXB_EQ_EULER3D ? h->cfg.nz : 1) * h->F : 0;
}
int64_t wxb_recv_size(const wxb_handle* h) {
    return h ? (int64_t)h->cfg.nghost * h->ntr * (h->cfg.equations == WXB_EQ_EULER3D ? h->cfg.nz : 1) * h->F : 0;
}
double* wxb_send_buffer(wxb_handle* h) { return h ? h->d_sendbuf : nullptr; }
double* wxb_recv_buffer(wxb_handle* h) { return h ? h->d_trace + h->trace_local : nullptr; }

int wxb_traces(wxb_handle* h, const double* q, const double* v, double eps, void* stream) {
    if (!h) return WXB_EINVAL;
    if (!q) {
        set_err(h, "null state pointer");
        return WXB_EINVAL;
    }
    return do_traces(h, q, v, eps, (cudaStream_t)stream);
}

int wxb_finish(wxb_handle* h, const double* q, const double* v, const double* rhs0, double dt, double eps, int mode,
               int part, double* out, void* stream) {
    if (!h) return WXB_EINVAL;
    int r = check_mode(h, q, v, rhs0, eps, mode, out);
    if (r != WXB_OK) return r;
    if (!h->traces_done) {
        set_err(h, "wxb_finish called before wxb_traces");
        return WXB_ESTATE;
    }
    r = do_finish(h, q, v, rhs0, dt, eps, mode, part, out, (cudaStream_t)stream);
    if (part != 1) h->traces_done = false;
    return r;
}

int wxb_rhs(wxb_handle* h, const double* q, double* dq, void* stream) {
    if (!h) return WXB_EINVAL;
    int r = check_mode(h, q, nullptr, nullptr, 0.0, -1, dq);
    if (r != WXB_OK) return r;
    if (h->cfg.nghost != 0) {
        set_err(h, "handle has remote neighbours: use wxb_traces / exchange / wxb_finish");
        return WXB_ESTATE;
    }
    r = do_traces(h, q, nullptr, 0.0, (cudaStream_t)stream);
    if (r != WXB_OK) return r;
    h->traces_done = false;
    return do_finish(h, q, nullptr, nullptr, 0.0, 0.0, -1, 0, dq, (cudaStream_t)stream);
}

int wxb_jvp(wxb_handle* h, const double* q, const double* v, const double* rhs0, double dt, double eps, int mode,
            double* out, void* stream) {
    if (!h) return WXB_EINVAL;
    if (mode < 0) {
        set_err(h, "wxb_jvp needs mode WXB_JVP_FD or WXB_JVP_RAT");
        return WXB_EINVAL;
    }
    int r = check_mode(h, q, v, rhs0, eps, mode, out);
    if (r != WXB_OK) return r;
    if (h->cfg.nghost != 0) {
        set_err(h, "handle has remote neighbours: use wxb_traces / exchange / wxb_finish");
        return WXB_ESTATE;
    }
    r = do_traces(h, q, v, eps, (cudaStream_t)stream);
    if (r != WXB_OK) return r;
    h->traces_done = false;
    return do_finish(h, q, v, rhs0, dt, eps, mode, 0, out, (cudaStream_t)stream);
}

}  // extern "C"

// ---- halo exchange inside the library -------------------------------------------------------------------------
// NCCL is reached through dlopen (the copy the process has already loaded, e.g. torch's, else the system one): the
// library carries no link-time dependency on it and single-GPU users never touch it.
struct WxbNcclId {
    char internal[WXB_NCCL_ID_BYTES];
};
namespace {
struct NcclApi {
    void* lib;
    int (*GetUniqueId)(void*);
    int (*CommInitRank)(void**, int, WxbNcclId, int);
    int (*CommDestroy)(void*);
    int (*GroupStart)();
    int (*GroupEnd)();
    int (*Send)(const void*, size_t, int, int, void*, cudaStream_t);
    int (*Recv)(void*, size_t, int, int, void*, cudaStream_t);
    const char* (*GetErrorString)(int);
};
NcclApi g_nccl = {};

int load_nccl(wxb_handle* h) {
    if (g_nccl.lib) return WXB_OK;
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) {
        set_err(h, "cannot load libnccl.so.2: %s", dlerror());
        return WXB_ECUDA;
    }
    NcclApi a = {};
    a.lib = lib;
    *(void**)&a.GetUniqueId = dlsym(lib, "ncclGetUniqueId");
    *(void**)&a.CommInitRank = dlsym(lib, "ncclCommInitRank");
    *(void**)&a.CommDestroy = dlsym(lib, "ncclCommDestroy");
    *(void**)&a.GroupStart = dlsym(lib, "ncclGroupStart");
    *(void**)&a.GroupEnd = dlsym(lib, "ncclGroupEnd");
    *(void**)&a.Send = dlsym(lib, "ncclSend");
    *(void**)&a.Recv = dlsym(lib, "ncclRecv");
    *(void**)&a.GetErrorString = dlsym(lib, "ncclGetErrorString");
    if (!a.GetUniqueId || !a.CommInitRank || !a.CommDestroy || !a.GroupStart || !a.GroupEnd || !a.Send || !a.Recv) {
        set_err(h, "libnccl.so.2 lacks a required symbol");
        return WXB_ECUDA;
    }
    g_nccl = a;
    return WXB_OK;
}

#define NC(call)                                                                                                   \
    do {                                                                                                           \
        int _e = (call);                                                                                           \
        if (_e != 0) {                                                                                             \
            set_err(h, "%s failed: %s", #call, g_nccl.GetErrorString ? g_nccl.GetErrorString(_e) : "NCCL error"); \
            return WXB_ECUDA;                                                                                      \
        }                                                                                                          \
    } while (0)

void comm_free(wxb_handle* h) {
    if (h->nccl_comm && g_nccl.CommDestroy) g_nccl.CommDestroy(h->nccl_comm);
    h->nccl_comm = nullptr;
    if (h->s_comm) cudaStreamDestroy(h->s_comm);
    if (h->ev_packed) cudaEventDestroy(h->ev_packed);
    if (h->ev_halo) cudaEventDestroy(h->ev_halo);
    h->s_comm = nullptr;
    h->ev_packed = h->ev_halo = nullptr;
    for (int** p : {&h->send_peer, &h->send_first, &h->send_count, &h->recv_peer, &h->recv_first, &h->recv_count}) {
        free(*p);
        *p = nullptr;
    }
}

int copy_list(int** dst, const int32_t* src, int n) {
    *dst = (int*)malloc(sizeof(int) * (n > 0 ? n : 1));
    if (!*dst) return WXB_ENOMEM;
    for (int i = 0; i < n; ++i) (*dst)[i] = src[i];
    return WXB_OK;
}
}  // namespace

extern "C" {

int wxb_comm_unique_id(void* id_out) {
    wxb_handle* h = nullptr;
    if (!id_out) return WXB_EINVAL;
    int r = load_nccl(h);
    if (r != WXB_OK) return r;
    NC(g_nccl.GetUniqueId(id_out));
    return WXB_OK;
}

int wxb_comm_init(wxb_handle* h, const void* id, int rank, int nranks, int n_send_peers, const int32_t* send_peer,
                  const int32_t* send_first, const int32_t* send_count, int n_recv_peers, const int32_t* recv_peer,
                  const int32_t* recv_first, const int32_t* recv_count) {
    if (!h) return WXB_EINVAL;
    if (!id || rank < 0 || rank >= nranks || n_send_peers < 0 || n_recv_peers < 0) {
        set_err(h, "bad communicator arguments (rank %d of %d)", rank, nranks);
        return WXB_EINVAL;
    }
    long long ns = 0, nr = 0;
    for (int i = 0; i < n_send_peers; ++i) {
        if (send_peer[i] < 0 || send_peer[i] >= nranks || send_peer[i] == rank || send_first[i] != ns || send_count[i] <= 0) {
            set_err(h, "send list entry %d is not a contiguous slice of the send buffer for another rank", i);
            return WXB_EINVAL;
        }
        ns += send_count[i];
    }
    for (int i = 0; i < n_recv_peers; ++i) {
        if (recv_peer[i] < 0 || recv_peer[i] >= nranks || recv_peer[i] == rank || recv_first[i] != nr || recv_count[i] <= 0) {
            set_err(h, "receive list entry %d is not a contiguous slice of the ghost records from another rank", i);
            return WXB_EINVAL;
        }
        nr += recv_count[i];
    }
    if (ns != h->cfg.nsend || nr != h->cfg.nghost) {
        set_err(h, "peer lists cover %lld send / %lld ghost records, the handle has %d / %d", ns, nr, h->cfg.nsend, h->cfg.nghost);
        return WXB_EINVAL;
    }
    int r = load_nccl(h);
    if (r != WXB_OK) return r;
    CU(cudaSetDevice(h->device));
    comm_free(h);
    WxbNcclId uid;
    memcpy(uid.internal, id, WXB_NCCL_ID_BYTES);
    NC(g_nccl.CommInitRank(&h->nccl_comm, nranks, uid, rank));
    CU(cudaStreamCreateWithFlags(&h->s_comm, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&h->ev_packed, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&h->ev_halo, cudaEventDisableTiming));
    h->comm_rank = rank;
    h->comm_nranks = nranks;
    h->n_send_peers = n_send_peers;
    h->n_recv_peers = n_recv_peers;
    if ((r = copy_list(&h->send_peer, send_peer, n_send_peers)) || (r = copy_list(&h->send_first, send_first, n_send_peers)) ||
        (r = copy_list(&h->send_count, send_count, n_send_peers)) || (r = copy_list(&h->recv_peer, recv_peer, n_recv_peers)) ||
        (r = copy_list(&h->recv_first, recv_first, n_recv_peers)) || (r = copy_list(&h->recv_count, recv_count, n_recv_peers)))
        return r;
    return WXB_OK;
}

int wxb_exchange(wxb_handle* h, void* stream) {
    if (!h) return WXB_EINVAL;
    if (!h->nccl_comm) {
        set_err(h, "wxb_exchange before wxb_comm_init");
        return WXB_ESTATE;
    }
    CU(cudaSetDevice(h->device));
    const size_t rec = (size_t)h->ntr * (h->cfg.equations == WXB_EQ_EULER3D ? h->cfg.nz : 1) * h->F;  // doubles per record
    double* recv = h->d_trace + h->trace_local;
    NC(g_nccl.GroupStart());
    for (int i = 0; i < h->n_send_peers; ++i)
        NC(g_nccl.Send(h->d_sendbuf + (size_t)h->send_first[i] * rec, (size_t)h->send_count[i] * rec, 8 /* ncclFloat64 */, h->send_peer[i],
                       h->nccl_comm, (cudaStream_t)stream));
    for (int i = 0; i < h->n_recv_peers; ++i)
        NC(g_nccl.Recv(recv + (size_t)h->recv_first[i] * rec, (size_t)h->recv_count[i] * rec, 8, h->recv_peer[i], h->nccl_comm, (cudaStream_t)stream));
    NC(g_nccl.GroupEnd());
    return WXB_OK;
}

// One whole evaluation on a handle WITH remote neighbours: boundary traces -> exchange on the handle's communication
// stream, overlapped with the tiles / elements that read no ghost record -> the rest.  mode -1 = plain RHS.
int wxb_eval_dist(wxb_handle* h, const double* q, const double* v, const double* rhs0, double dt, double eps, int mode, double* out,
                  void* stream) {
    if (!h) return WXB_EINVAL;
    int r = check_mode(h, q, mode >= 0 ? v : nullptr, mode >= 0 ? rhs0 : nullptr, eps, mode, out);
    if (r != WXB_OK) return r;
    if (h->cfg.nghost != 0 && !h->nccl_comm) {
        set_err(h, "handle has remote neighbours: call wxb_comm_init first (or use wxb_traces / your own exchange / wxb_finish)");
        return WXB_ESTATE;
    }
    cudaStream_t s = (cudaStream_t)stream;
    const double* vv = mode >= 0 ? v : nullptr;
    r = do_traces(h, q, vv, eps, s);
    if (r != WXB_OK) return r;
    if (h->cfg.nghost != 0) {
        CU(cudaEventRecord(h->ev_packed, s));
        CU(cudaStreamWaitEvent(h->s_comm, h->ev_packed, 0));
        r = wxb_exchange(h, h->s_comm);
        if (r != WXB_OK) return r;
        CU(cudaEventRecord(h->ev_halo, h->s_comm));
        r = do_finish(h, q, vv, rhs0, dt, eps, mode, 1, out, s);
        if (r != WXB_OK) return r;
        CU(cudaStreamWaitEvent(s, h->ev_halo, 0));
        r = do_finish(h, q, vv, rhs0, dt, eps, mode, 2, out, s);
    } else {
        r = do_finish(h, q, vv, rhs0, dt, eps, mode, 0, out, s);
    }
    h->traces_done = false;
    return r;
}

static int host_stage(wxb_handle* h) {
    const size_t bytes = (size_t)wxb_state_size(h) * sizeof(double);
    if (!h->d_q) {
        CU(cudaSetDevice(h->device));
        CU(cudaMalloc((void**)&h->d_q, bytes));
        CU(cudaMalloc((void**)&h->d_v, bytes));
        CU(cudaMalloc((void**)&h->d_r0, bytes));
        CU(cudaMalloc((void**)&h->d_out, bytes));
    }
    return WXB_OK;
}

// ---- chunked host pipeline ----------------------------------------------------------------------------------
// The host entry points move a whole state each way over PCIe; done back to back (upload, kernels, download) the
// link idles in one direction at a time.  For the 3-D Euler equations the state is cut into chunks of vertical
// levels: horizontal neighbours live on the same level, vertical ones at most one level away, so the main kernel of
// chunk c only needs the traces of chunks <= c+1.  Upload of chunk c+2, kernels of chunk c and download of chunk
// c-1 then run concurrently on three streams.  Shallow water (one level) degenerates to one chunk.
static int pipe_setup(wxb_handle* h) {
    if (h->pipe_ready) return WXB_OK;
    int r = host_stage(h);
    if (r != WXB_OK) return r;
    const bool euler = h->cfg.equations == WXB_EQ_EULER3D;
    const int nz = euler ? h->cfg.nz : 1;
    // measured at S2 s = 1 (24 levels), end-to-end step: 8 chunks 172 ms, 12 chunks 165 ms, 24 chunks 158 ms
    int want = getenv("WXB_HOST_CHUNKS") ? atoi(getenv("WXB_HOST_CHUNKS")) : 24;
    if (want < 1) want = 1;
    if (want > WXB_MAX_CHUNKS) want = WXB_MAX_CHUNKS;
    h->nchunk = nz < want ? nz : want;
    std::vector<int> list((size_t)h->E);
    int pos = 0;
    for (int c = 0; c < h->nchunk; ++c) {
        const int k0 = (int)((long long)c * nz / h->nchunk), k1 = (int)((long long)(c + 1) * nz / h->nchunk);
        h->chunk_k[c] = k0;
        h->chunk_off[c] = pos;
        for (int s = 0; s < h->st.nslabs; ++s) {
            const int per_level = h->st.ny[s] * h->cfg.nx;
            for (int el = k0 * per_level; el < k1 * per_level; ++el) list[pos++] = h->st.ebase[s] + el;
        }
    }
    h->chunk_k[h->nchunk] = nz;
    h->chunk_off[h->nchunk] = pos;
    if (pos != h->E) {
        set_err(h, "internal: chunk lists cover %d of %d elements", pos, h->E);
        return WXB_ESTATE;
    }
    UP(h->d_clist, list.data(), list.size());
    CU(cudaStreamCreate(&h->s_in));
    CU(cudaStreamCreate(&h->s_cmp));
    CU(cudaStreamCreate(&h->s_out));
    for (int c = 0; c < h->nchunk; ++c) {
        CU(cudaEventCreateWithFlags(&h->ev_in[c], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&h->ev_main[c], cudaEventDisableTiming));
    }
    if (h->cfg.nghost != 0 || h->cfg.nsend != 0) {
        for (int c = 0; c < h->nchunk; ++c) {
            CU(cudaEventCreateWithFlags(&h->ev_packed_c[c], cudaEventDisableTiming));
            CU(cudaEventCreateWithFlags(&h->ev_halo_c[c], cudaEventDisableTiming));
        }
        if (h->cfg.nghost > 0) CU(cudaMalloc((void**)&h->d_recv_stage, (size_t)h->cfg.nghost * h->ntr * nz * h->F * sizeof(double)));
    }
    h->pipe_ready = true;
    return WXB_OK;
}

// copy the levels of chunk c of one state-shaped array (per slab: nvar rows of contiguous levels)
static int chunk_copy(wxb_handle* h, int c, double* dst, const double* src, cudaMemcpyKind kind, cudaStream_t s) {
    if (h->nchunk == 1) {
        CU(cudaMemcpyAsync(dst, src, (size_t)wxb_state_size(h) * sizeof(double), kind, s));
        return WXB_OK;
    }
    const int nz = h->cfg.nz;
    const int k0 = h->chunk_k[c], k1 = h->chunk_k[c + 1];
    for (int sl = 0; sl < h->st.nslabs; ++sl) {
        const size_t level = (size_t)h->st.ny[sl] * h->cfg.nx * h->P;  // doubles per level of this slab
        const size_t off = (size_t)h->nvar * h->P * h->st.ebase[sl] + (size_t)k0 * level;
        const size_t pitch = (size_t)nz * level * sizeof(double);
        CU(cudaMemcpy2DAsync(dst + off, pitch, src + off, pitch, (size_t)(k1 - k0) * level * sizeof(double), h->nvar, kind, s));
    }
    return WXB_OK;
}

static bool has_remote(const wxb_handle* h) { return h->cfg.nghost != 0 || h->cfg.nsend != 0; }

// Multi-rank handle on a path the level pipeline does not cover (pair of kernels, shallow water): whole arrays up,
// wxb_eval_dist, whole array down -- still one library call, the exchange overlapped with the inner elements.
static int whole_array_dist_eval(wxb_handle* h, double* d_q, const double* q_host, double* d_v, const double* v_host, double* d_r0,
                                 const double* r0_host, double dt, double eps, int mode, double* d_out, double* out_host) {
    CU(cudaSetDevice(h->device));
    const size_t bytes = (size_t)wxb_state_size(h) * sizeof(double);
    cudaStream_t sc = h->s_cmp;
    if (q_host) CU(cudaMemcpyAsync(d_q, q_host, bytes, cudaMemcpyHostToDevice, sc));
    if (mode >= 0 && v_host) CU(cudaMemcpyAsync(d_v, v_host, bytes, cudaMemcpyHostToDevice, sc));
    if (mode >= 0 && r0_host) CU(cudaMemcpyAsync(d_r0, r0_host, bytes, cudaMemcpyHostToDevice, sc));
    int r = wxb_eval_dist(h, d_q, mode >= 0 ? d_v : nullptr, mode >= 0 ? d_r0 : nullptr, dt, eps, mode, d_out, sc);
    if (r != WXB_OK) return r;
    if (out_host) CU(cudaMemcpyAsync(out_host, d_out, bytes, cudaMemcpyDeviceToHost, sc));
    CU(cudaStreamSynchronize(sc));
    return WXB_OK;
}

// One evaluation with host operands.  *_host == nullptr means "already on the device" (in the matching d_* array).
// Multi-rank handles on the fused Euler path: the halo exchange is cut into the same level chunks.  As soon as chunk
// c has arrived, euler_edge_traces forms the records other ranks need of ITS levels (a record is level-local), NCCL
// moves them on the handle's communication stream and wxb_unpack_ghost files them under the ghost records; the sweep of
// chunk c-1 follows.  Uploads, exchange, sweeps and downloads of different chunks overlap as on one rank.
static int pipelined_eval(wxb_handle* h, double* d_q, const double* q_host, double* d_v, const double* v_host, double* d_r0,
                          const double* r0_host, double dt, double eps, int mode, double* d_out, double* out_host) {
    CU(cudaSetDevice(h->device));
    const bool euler = h->cfg.equations == WXB_EQ_EULER3D;
    const bool jvp = mode >= 0;
    int r = check_mode(h, d_q, jvp ? d_v : nullptr, jvp ? d_r0 : nullptr, eps, mode, d_out);
    if (r != WXB_OK) return r;
    const bool fused = use_fused(h, jvp);
    const bool dist = has_remote(h);
    // With more than two ranks copying at once the boxes this was measured on run out of HOST bandwidth (4 GPUs: 17-24
    // GB/s per GPU and direction when all copy both ways, against 45-55 one way): overlapping the two directions then
    // buys nothing and the pipeline's fill and drain cost ~20 %.  WXB_HOST_DIST=pipe|whole overrides.
    const char* hd_env = getenv("WXB_HOST_DIST");
    const bool dist_pipe = hd_env ? strcmp(hd_env, "pipe") == 0 : h->comm_nranks <= 2;
    if (dist && !(fused && euler && h->nchunk > 1 && dist_pipe))
        return whole_array_dist_eval(h, d_q, q_host, d_v, v_host, d_r0, r0_host, dt, eps, mode, d_out, out_host);
    if (jvp && !h->d_qp && !fused) CU(cudaMalloc((void**)&h->d_qp, (size_t)wxb_state_size(h) * sizeof(double)));
    const int nc = h->nchunk;
    // uploads, chunk by chunk
    for (int c = 0; c < nc; ++c) {
        if (q_host && (r = chunk_copy(h, c, d_q, q_host, cudaMemcpyHostToDevice, h->s_in)) != WXB_OK) return r;
        if (jvp && v_host && (r = chunk_copy(h, c, d_v, v_host, cudaMemcpyHostToDevice, h->s_in)) != WXB_OK) return r;
        if (jvp && r0_host && (r = chunk_copy(h, c, d_r0, r0_host, cudaMemcpyHostToDevice, h->s_in)) != WXB_OK) return r;
        CU(cudaEventRecord(h->ev_in[c], h->s_in));
    }
    EvalArgs at, am;
    fill_args(h, at, d_q, jvp ? d_v : nullptr, nullptr, 0.0, eps, -1, nullptr);
    fill_args(h, am, d_q, jvp ? d_v : nullptr, jvp ? d_r0 : nullptr, dt, eps, mode, d_out);
    h->ev = h->ev_ring[h->nevals % WXB_TIMING_RING];
    cudaStream_t sc = h->s_cmp;
    CU(cudaEventRecord(h->ev[0], sc));
    const size_t rec = (size_t)h->ntr * h->F;  // doubles per record and level
    // records of chunk c for the other ranks -> exchange -> ghost records of chunk c (stream s_comm)
    auto halo = [&](int c) -> int {
        const int k0 = h->chunk_k[c], k1 = h->chunk_k[c + 1];
        const size_t rc = rec * (size_t)(k1 - k0);  // doubles per record in this chunk
        at.k0 = k0;
        at.k1 = k1;
        at.sendbuf = h->d_sendbuf + (size_t)h->cfg.nsend * k0 * rec;
        if (wxb_euler_edge_traces(at, h->n, sc) != 0) {
            set_err(h, "trace kernel launch failed: %s", cudaGetErrorString(cudaGetLastError()));
            return WXB_ECUDA;
        }
        CU(cudaEventRecord(h->ev_packed_c[c], sc));
        CU(cudaStreamWaitEvent(h->s_comm, h->ev_packed_c[c], 0));
        double* stage = h->d_recv_stage + (size_t)h->cfg.nghost * k0 * rec;
        NC(g_nccl.GroupStart());
        for (int i = 0; i < h->n_send_peers; ++i)
            NC(g_nccl.Send(at.sendbuf + (size_t)h->send_first[i] * rc, (size_t)h->send_count[i] * rc, 8 /* ncclFloat64 */, h->send_peer[i],
                           h->nccl_comm, h->s_comm));
        for (int i = 0; i < h->n_recv_peers; ++i)
            NC(g_nccl.Recv(stage + (size_t)h->recv_first[i] * rc, (size_t)h->recv_count[i] * rc, 8, h->recv_peer[i], h->nccl_comm, h->s_comm));
        NC(g_nccl.GroupEnd());
        if (wxb_unpack_ghost(stage, h->d_trace + h->trace_local, h->cfg.nghost, (int)rec, k0, k1, h->cfg.nz, h->s_comm) != 0) {
            set_err(h, "unpack kernel launch failed: %s", cudaGetErrorString(cudaGetLastError()));
            return WXB_ECUDA;
        }
        CU(cudaEventRecord(h->ev_halo_c[c], h->s_comm));
        return WXB_OK;
    };
    auto traces = [&](int c) -> int {
        if (fused) return 0;  // the fused kernel extrapolates from the state itself
        if (nc > 1) {
            at.elist = h->d_clist + h->chunk_off[c];
            at.nelist = h->chunk_off[c + 1] - h->chunk_off[c];
        }
        return euler ? wxb_euler_traces(at, h->n, sc) : wxb_sw_traces(at, h->n, sc);
    };
    auto maink = [&](int c) -> int {
        if (fused && !euler) return wxb_sw_tile(am, h->n, sc);
        if (fused) {  // levels of chunk c; the levels either side of it have been uploaded (see the loop below)
            am.k0 = h->chunk_k[c];
            am.k1 = h->chunk_k[c + 1];
            return wxb_euler_sweep(am, h->n, sc);
        }
        if (nc > 1) {
            am.elist = h->d_clist + h->chunk_off[c];
            am.nelist = h->chunk_off[c + 1] - h->chunk_off[c];
        }
        return euler ? wxb_euler_main(am, h->n, sc) : wxb_sw_main(am, h->n, sc);
    };
    for (int c = 0; c <= nc; ++c) {
        if (c < nc) {
            CU(cudaStreamWaitEvent(sc, h->ev_in[c], 0));
            if (traces(c) != 0) {
                set_err(h, "trace kernel launch failed: %s", cudaGetErrorString(cudaGetLastError()));
                return WXB_ECUDA;
            }
            if (dist && (r = halo(c)) != WXB_OK) return r;
            if (c == nc - 1) CU(cudaEventRecord(h->ev[1], sc));
        }
        if (c >= 1) {  // traces of chunks <= c are in flight: chunk c-1 has everything it reads
            if (dist) CU(cudaStreamWaitEvent(sc, h->ev_halo_c[c < nc ? c : nc - 1], 0));  // ghost records of chunks <= c are in place
            if (maink(c - 1) != 0) {
                set_err(h, "main kernel launch failed: %s", cudaGetErrorString(cudaGetLastError()));
                return WXB_ECUDA;
            }
            CU(cudaEventRecord(h->ev_main[c - 1], sc));
            CU(cudaStreamWaitEvent(h->s_out, h->ev_main[c - 1], 0));
            if (out_host && (r = chunk_copy(h, c - 1, out_host, d_out, cudaMemcpyDeviceToHost, h->s_out)) != WXB_OK) return r;
        }
    }
    CU(cudaEventRecord(h->ev[2], sc));
    h->ev_mode[h->nevals % WXB_TIMING_RING] = mode;
    h->nevals++;
    h->timed = true;
    h->traces_done = false;
    CU(cudaStreamSynchronize(h->s_out));
    CU(cudaStreamSynchronize(sc));
    return WXB_OK;
}

// host entry points on a handle with remote neighbours need the library's own exchange (wxb_comm_init) and are then
// collective like wxb_eval_dist; the advection-only short cut keeps the plain path
static int host_ready(wxb_handle* h, bool* plain) {
    if (has_remote(h) && !h->nccl_comm) {
        set_err(h, "handle has remote neighbours: call wxb_comm_init first, or use wxb_traces / exchange / wxb_finish");
        return WXB_ESTATE;
    }
    *plain = h->cfg.equations == WXB_EQ_EULER3D && h->cfg.advection_only;
    return *plain ? host_stage(h) : pipe_setup(h);
}

int wxb_rhs_host(wxb_handle* h, const double* q_host, double* dq_host) {
    if (!h) return WXB_EINVAL;
    if (!q_host || !dq_host) {
        set_err(h, "null host pointer");
        return WXB_EINVAL;
    }
    bool plain;
    int r = host_ready(h, &plain);
    if (r != WXB_OK) return r;
    if (!plain) return pipelined_eval(h, h->d_q, q_host, nullptr, nullptr, nullptr, nullptr, 0.0, 0.0, -1, h->d_out, dq_host);
    const size_t bytes = (size_t)wxb_state_size(h) * sizeof(double);
    CU(cudaMemcpyAsync(h->d_q, q_host, bytes, cudaMemcpyHostToDevice, 0));
    r = wxb_rhs(h, h->d_q, h->d_out, nullptr);
    if (r != WXB_OK) return r;
    CU(cudaMemcpyAsync(dq_host, h->d_out, bytes, cudaMemcpyDeviceToHost, 0));
    CU(cudaStreamSynchronize(0));
    return WXB_OK;
}

int wxb_jvp_host(wxb_handle* h, const double* q_host, const double* v_host, const double* rhs0_host, double dt,
                 double eps, int mode, double* out_host) {
    if (!h) return WXB_EINVAL;
    if (!q_host || !v_host || !rhs0_host || !out_host) {
        set_err(h, "null host pointer");
        return WXB_EINVAL;
    }
    if (mode < 0) {
        set_err(h, "wxb_jvp_host needs mode WXB_JVP_FD or WXB_JVP_RAT");
        return WXB_EINVAL;
    }
    bool plain;
    int r = host_ready(h, &plain);
    if (r != WXB_OK) return r;
    if (!plain) return pipelined_eval(h, h->d_q, q_host, h->d_v, v_host, h->d_r0, rhs0_host, dt, eps, mode, h->d_out, out_host);
    const size_t bytes = (size_t)wxb_state_size(h) * sizeof(double);
    CU(cudaMemcpyAsync(h->d_q, q_host, bytes, cudaMemcpyHostToDevice, 0));
    CU(cudaMemcpyAsync(h->d_v, v_host, bytes, cudaMemcpyHostToDevice, 0));
    CU(cudaMemcpyAsync(h->d_r0, rhs0_host, bytes, cudaMemcpyHostToDevice, 0));
    r = wxb_jvp(h, h->d_q, h->d_v, h->d_r0, dt, eps, mode, h->d_out, nullptr);
    if (r != WXB_OK) return r;
    CU(cudaMemcpyAsync(out_host, h->d_out, bytes, cudaMemcpyDeviceToHost, 0));
    CU(cudaStreamSynchronize(0));
    return WXB_OK;
}

static int lin_stage(wxb_handle* h) {
    const size_t bytes = (size_t)wxb_state_size(h) * sizeof(double);
    if (!h->d_lq) {
        CU(cudaSetDevice(h->device));
        CU(cudaMalloc((void**)&h->d_lq, bytes));
        CU(cudaMalloc((void**)&h->d_lr, bytes));
    }
    return WXB_OK;
}

int wxb_linearize_host(wxb_handle* h, const double* q_host, double* rhs_host) {
    if (!h) return WXB_EINVAL;
    if (!q_host || !rhs_host) {
        set_err(h, "null host pointer");
        return WXB_EINVAL;
    }
    bool plain;
    int r = host_ready(h, &plain);
    if (r == WXB_OK) r = lin_stage(h);
    if (r != WXB_OK) return r;
    h->bound = false;
    if (!plain) {
        r = pipelined_eval(h, h->d_lq, q_host, nullptr, nullptr, nullptr, nullptr, 0.0, 0.0, -1, h->d_lr, rhs_host);
    } else {
        const size_t bytes = (size_t)wxb_state_size(h) * sizeof(double);
        CU(cudaMemcpyAsync(h->d_lq, q_host, bytes, cudaMemcpyHostToDevice, 0));
        r = wxb_rhs(h, h->d_lq, h->d_lr, nullptr);
        if (r != WXB_OK) return r;
        CU(cudaMemcpyAsync(rhs_host, h->d_lr, bytes, cudaMemcpyDeviceToHost, 0));
        CU(cudaStreamSynchronize(0));
    }
    h->bound = r == WXB_OK;
    return r;
}

int wxb_bind_host(wxb_handle* h, const double* q_host, const double* rhs0_host) {
    if (!h) return WXB_EINVAL;
    if (!q_host || !rhs0_host) {
        set_err(h, "null host pointer");
        return WXB_EINVAL;
    }
    bool plain;
    int r = host_ready(h, &plain);
    if (r == WXB_OK) r = lin_stage(h);
    if (r != WXB_OK) return r;
    h->bound = false;
    const size_t bytes = (size_t)wxb_state_size(h) * sizeof(double);
    CU(cudaMemcpyAsync(h->d_lq, q_host, bytes, cudaMemcpyHostToDevice, 0));
    CU(cudaMemcpyAsync(h->d_lr, rhs0_host, bytes, cudaMemcpyHostToDevice, 0));
    CU(cudaStreamSynchronize(0));
    h->bound = true;
    return WXB_OK;
}

int wxb_jvp_bound_host(wxb_handle* h, const double* v_host, double dt, double eps, int mode, double* out_host) {
    if (!h) return WXB_EINVAL;
    if (!v_host || !out_host) {
        set_err(h, "null host pointer");
        return WXB_EINVAL;
    }
    if (mode < 0) {
        set_err(h, "wxb_jvp_bound_host needs mode WXB_JVP_FD or WXB_JVP_RAT");
        return WXB_EINVAL;
    }
    if (!h->bound) {
        set_err(h, "no linearisation point: call wxb_linearize_host or wxb_bind_host first");
        return WXB_ESTATE;
    }
    bool plain;
    int r = host_ready(h, &plain);
    if (r != WXB_OK) return r;
    if (!plain) return pipelined_eval(h, h->d_lq, nullptr, h->d_v, v_host, h->d_lr, nullptr, dt, eps, mode, h->d_out, out_host);
    const size_t bytes = (size_t)wxb_state_size(h) * sizeof(double);
    CU(cudaMemcpyAsync(h->d_v, v_host, bytes, cudaMemcpyHostToDevice, 0));
    r = wxb_jvp(h, h->d_lq, h->d_v, h->d_lr, dt, eps, mode, h->d_out, nullptr);
    if (r != WXB_OK) return r;
    CU(cudaMemcpyAsync(out_host, h->d_out, bytes, cudaMemcpyDeviceToHost, 0));
    CU(cudaStreamSynchronize(0));
    return WXB_OK;
}

int wxb_timings(wxb_handle* h, float out_ms[3]) {
    if (!h || !out_ms) return WXB_EINVAL;
    if (!h->timed) {
        set_err(h, "no completed evaluation to time");
        return WXB_ESTATE;
    }
    CU(cudaSetDevice(h->device));
    cudaEvent_t* ev = h->ev_ring[(h->nevals - 1) % WXB_TIMING_RING];
    CU(cudaEventSynchronize(ev[2]));
    CU(cudaEventElapsedTime(&out_ms[0], ev[0], ev[1]));
    CU(cudaEventElapsedTime(&out_ms[1], ev[1], ev[2]));
    CU(cudaEventElapsedTime(&out_ms[2], ev[0], ev[2]));
    return WXB_OK;
}

int wxb_timing_history(wxb_handle* h, int max_evals, float* out_ms, int* out_mode) {
    if (!h || !out_ms || max_evals < 0) return WXB_EINVAL;
    CU(cudaSetDevice(h->device));
    long long n = h->nevals < max_evals ? h->nevals : max_evals;
    if (n > WXB_TIMING_RING) n = WXB_TIMING_RING;
    for (long long i = 0; i < n; ++i) {
        const long long id = h->nevals - n + i;  // oldest first
        cudaEvent_t* ev = h->ev_ring[id % WXB_TIMING_RING];
        CU(cudaEventSynchronize(ev[2]));
        CU(cudaEventElapsedTime(&out_ms[3 * i + 0], ev[0], ev[1]));
        CU(cudaEventElapsedTime(&out_ms[3 * i + 1], ev[1], ev[2]));
        CU(cudaEventElapsedTime(&out_ms[3 * i + 2], ev[0], ev[2]));
        if (out_mode) out_mode[i] = h->ev_mode[id % WXB_TIMING_RING];
    }
    return (int)n;
}

}  // extern "C"
