// Internal declarations shared by the kernels and the C ABI (not part of the public interface).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/wxb200.h"

// wx_factory/common/definitions.py:5-12
#define WXB_P0 100000.0
#define WXB_RD 287.05
#define WXB_CPD 1005.46
#define WXB_CVD (WXB_CPD - WXB_RD)
#define WXB_GRAVITY 9.80616
#define WXB_GAMMA (WXB_CPD / WXB_CVD)
#define WXB_LOG_RD_P0 (-5.853269048356583)  // log(Rd / p0), correctly rounded
#define WXB_LOG_P0 11.512925464970229      // log(p0)
#define WXB_SQRT_GAMMA_P0 374.10732733784823  // sqrt((cpd/cvd) * p0), correctly rounded

#define WXB_TIMING_RING 256

struct SlabTab {
    int nslabs;
    int ny[WXB_MAX_SLABS];
    int ebase[WXB_MAX_SLABS + 1];     // first local element of each slab
    int cbase[WXB_MAX_SLABS + 1];     // first local column of each slab
    long long fmi[WXB_MAX_SLABS];     // offset (doubles) of the slab's block in fm_i / fm_j / fm_k
    long long fmj[WXB_MAX_SLABS];
    long long fmk[WXB_MAX_SLABS];
};

// Arguments of one evaluation (passed by value to every kernel).
struct EvalArgs {
    const double* q;      // state
    const double* v;      // perturbation direction or nullptr
    double eps;           // q + eps*v
    double inv_eps;       // 1/eps (0 when there is no direction): the JVP epilogue multiplies instead of dividing
    const double* rhs0;   // cached base RHS (JVP modes)
    double dt;
    int mode;             // -1 plain RHS, WXB_JVP_FD, WXB_JVP_RAT
    double* out;
    double* qp;           // scratch: q + eps*v (written by the traces kernel in JVP mode, read by the main kernel)
    double* trace;        // SW: [3][4][E][F] then ghost records [rec][3][1][F]; Euler: [E][6][6][F] then [rec][nz][6][F]
    double* sendbuf;
    const double* sqrtg;
    const double* h;
    const double* chris;
    const double* idz;
    const double* damp;
    const double* topo;
    const double* fmi;
    const double* fmj;
    const double* fmk;
    const double* topo_itf;
    const int4* nbr;
    const double* xform;
    const int* send;
    const double* ops;    // ext_w[n] ext_e[n] diff[n*n] corr[2n] hfilt[n*n] (Euler: + dlogp[n*n] klogp[n], wxb_api.cu)
    const int* elist;     // optional list of elements to process (nullptr = all)
    int nelist;
    // column-sweep path (euler_sweep.cu)
    const double* fasttab;  // log / exp tables (fastmath.cuh)
    const int4* tinfo;      // per tile: slab, first row, first column of the tile (built at create)
    const int* tlist;       // optional list of tiles to process (nullptr = all)
    int ntiles;             // blocks to launch
    int k0, k1;             // levels [k0, k1) to sweep
    int E, C, nz, nx, nsend;
    int has_damp, has_topo;
    double tau0;
    SlabTab st;
};

#define WXB_MAX_CHUNKS 64

struct wxb_handle {
    wxb_config cfg;
    int device;
    int n, P, F, nvar, ntr, nfaces;
    int E, C;
    SlabTab st;
    // device arrays
    double* d_metric;  // Euler: owner of d_chris / d_sqrtg / d_h / d_idz (one block, see wxb_api.cu)
    double *d_ops, *d_sqrtg, *d_h, *d_chris, *d_idz, *d_damp, *d_topo, *d_fmi, *d_fmj, *d_fmk, *d_topo_itf, *d_xform;
    double* d_expf;    // (n x n) nodal exponential filter (wxb_set_expfilter), stepops.cu
    int4* d_nbr;
    int *d_send, *d_elist_inner, *d_elist_outer;
    int n_inner, n_outer;
    // column-sweep path of the Euler equations (euler_sweep.cu): one launch per evaluation, no local trace array
    bool sweep;              // the fused path exists on this handle (euler_sweep.cu / sw_tile.cu)
    bool split;              // the two-kernel path exists (euler3d.cu / sw2d.cu: trace array, element lists)
    double* d_fasttab;
    double* d_trace_alloc;   // what was really allocated for d_trace (sweep: the ghost records only)
    int *d_tiles_inner, *d_tiles_outer;
    int4* d_tinfo;
    int nt_all, nt_inner, nt_outer;
    double *d_trace, *d_sendbuf, *d_qp;
    int64_t trace_local;  // doubles in the local part of the trace array
    // staging for the *_host entry points
    double *d_q, *d_v, *d_r0, *d_out;
    // linearisation point kept on the device between wxb_linearize_host / wxb_bind_host and wxb_jvp_bound_host
    double *d_lq, *d_lr;
    bool bound;
    // chunked host pipeline (k-level chunks: upload of chunk c+1, kernels of chunk c, download of chunk c-1 overlap)
    bool pipe_ready;
    int nchunk;
    int chunk_k[WXB_MAX_CHUNKS + 1];    // first level of every chunk
    int chunk_off[WXB_MAX_CHUNKS + 1];  // offsets into d_clist
    int* d_clist;                       // elements of chunk 0, chunk 1, ...
    cudaStream_t s_in, s_cmp, s_out;
    cudaEvent_t ev_in[WXB_MAX_CHUNKS], ev_main[WXB_MAX_CHUNKS];
    // multi-rank handles: the exchange runs chunk by chunk too (records of chunk c -> NCCL -> unpacked ghost records)
    cudaEvent_t ev_packed_c[WXB_MAX_CHUNKS], ev_halo_c[WXB_MAX_CHUNKS];
    double* d_recv_stage;       // ghost records as they arrive, chunk-major
    // one event triple per evaluation, in a ring: [0] start, [1] traces(+pack) done, [2] main done
    cudaEvent_t ev_ring[WXB_TIMING_RING][3];
    int ev_mode[WXB_TIMING_RING];
    cudaEvent_t* ev;      // triple of the evaluation in flight
    long long nevals;     // completed evaluations
    bool timed;
    bool traces_done;
    // halo exchange inside the library (wxb_comm_init / wxb_exchange): NCCL point-to-point on the handle's own stream
    void* nccl_comm;
    cudaStream_t s_comm;
    cudaEvent_t ev_packed, ev_halo;
    int comm_rank, comm_nranks;
    int n_send_peers, n_recv_peers;
    int *send_peer, *send_first, *send_count, *recv_peer, *recv_first, *recv_count;  // host arrays
    char err[512];
};

// launchers implemented per equation set
int wxb_euler_configure(int n);
int wxb_euler_traces(const EvalArgs& a, int n, cudaStream_t s);
int wxb_euler_main(const EvalArgs& a, int n, cudaStream_t s);
int wxb_sweep_configure(int n);
void wxb_sweep_tile_shape(int n, int* tj, int* ti);
int wxb_euler_sweep(const EvalArgs& a, int n, cudaStream_t s);
int wxb_euler_edge_traces(const EvalArgs& a, int n, cudaStream_t s);
int wxb_unpack_ghost(const double* stage, double* ghost, int nghost, int rec, int k0, int k1, int nz, cudaStream_t s);
void wxb_sw_tile_shape(int n, int* tj, int* ti);
int wxb_sw_tile(const EvalArgs& a, int n, cudaStream_t s);
int wxb_sw_tile_edge_traces(const EvalArgs& a, int n, cudaStream_t s);
int wxb_sw_traces(const EvalArgs& a, int n, cudaStream_t s);
int wxb_sw_main(const EvalArgs& a, int n, cudaStream_t s);
int wxb_pack(const EvalArgs& a, int ntr, int F, int nfaces, int elem_major, cudaStream_t s);

// decode a local element index into (slab, level k, local row jl, col i)
struct ElemPos {
    int s, k, jl, i, c;       // c = local column
    long long qoff;           // offset of (var 0, this element, node 0) in the state
    long long vstride;        // distance between variables in the state
    int ny;
};

__device__ __forceinline__ ElemPos decode_elem(const SlabTab& st, int e, int nz, int nx, int P, int nvar) {
    ElemPos r;
    int s = 0;
#pragma unroll
    for (int t = 1; t < WXB_MAX_SLABS; ++t)
        if (t < st.nslabs && e >= st.ebase[t]) s = t;
    const int ny = st.ny[s];
    const int el = e - st.ebase[s];
    r.s = s;
    r.ny = ny;
    r.i = el % nx;
    const int t2 = el / nx;
    r.jl = t2 % ny;
    r.k = t2 / ny;
    r.c = st.cbase[s] + r.jl * nx + r.i;
    r.vstride = (long long)nz * ny * nx * P;
    r.qoff = (long long)nvar * P * st.ebase[s] + (long long)el * P;
    return r;
}
