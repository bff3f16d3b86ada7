/* wxb200.h -- C ABI of the B200-native cubed-sphere DFR right-hand side and FD Jacobian-vector product.
 *
 * This is the drop-in boundary for WxFactory's hot path.  The reference has no FFI for this path (it is
 * numpy); the entry points below are what a binding for it would call, each one citing the reference
 * interface it replaces (paths relative to the reference root):
 *
 *   wxb_create        <- RHS.__init__ / RhsShallowWater.__init__ receiving geometry, operators, metric
 *                        (wx_factory/rhs/rhs.py:14-35, wx_factory/rhs/rhs_sw.py:18-36) and the halo tables of
 *                        ProcessTopology.__init__ (wx_factory/process_topology.py:50-267)
 *   wxb_rhs           <- RHS.__call__(q) -> dq/dt  (wx_factory/rhs/rhs.py:75-122),
 *                        RhsShallowWater.__call__   (wx_factory/rhs/rhs_sw.py:38-56)
 *   wxb_jvp           <- matvec_fun(vec, dt, Q, rhs, rhs_handle, "fd") (wx_factory/solvers/matvec.py:36-68)
 *                        and matvec_rat (wx_factory/solvers/matvec.py:76-88)
 *   wxb_traces / wxb_finish
 *                     <- the two halves of RHS.__call__ either side of the halo exchange:
 *                        solution_extrapolation + start_communication (wx_factory/rhs/rhs_dfr.py:50-71,141-172)
 *                        and end_communication ... forcing_terms     (wx_factory/rhs/rhs_dfr.py:174-313)
 *   wxb_rhs_host / wxb_jvp_host
 *                     <- the same calls with HOST buffers (what the reference's CPU integrators hold)
 *   wxb_linearize_host / wxb_bind_host / wxb_jvp_bound_host
 *                     <- MatvecOp.__init__ capturing (dt, Q, rhs) and MatvecOp.__call__(vec)
 *                        (wx_factory/solvers/matvec.py:7-34,71-73; wx_factory/integrators/ros2.py:29)
 *   wxb_timings       <- RHS.retrieve_last_times / Device.elapsed (wx_factory/rhs/rhs.py:72-73,
 *                        wx_factory/device/device.py:201-214)
 *   wxb_destroy, wxb_last_error, wxb_version : lifetime / error convention (the reference raises Python
 *                        exceptions: wx_factory/rhs/rhs_selector.py:44,83)
 *   wxb_set_expfilter / wxb_expfilter
 *                     <- DFROperators.apply_filter_3d (wx_factory/geometry/operators.py:101-119,257-261)
 *   wxb_vec_*         <- the full-state vector algebra of the callers: Tvdrk3 / Ros2 stage combinations
 *                        (wx_factory/integrators/tvdrk3.py:13-17, ros2.py:30), the Gram-Schmidt products and updates of
 *                        FGMRES, KIOPS and PMEX (wx_factory/solvers/fgmres.py:16-72, kiops.py:176-211, pmex.py:150-230),
 *                        the NaN check (wx_factory/simulation/simulation.py:399-408); declared in the second half
 *
 * Conventions: plain pointers and sizes only; every function returns 0 on success or a negative WXB_E*
 * code and never throws; `stream` is a cudaStream_t passed as void* (NULL = legacy default stream); work is
 * enqueued asynchronously on `stream` (the *_host variants synchronise before returning).  A handle is not
 * re-entrant: calls on one handle must be serialised by the caller, as in the reference
 * (wx_factory/rhs/rhs_dfr.py:148-172 keeps request state between start/end_communication).
 *
 * State layout (device or host, float64, C order): the slabs this handle owns, concatenated; each slab is
 * the reference's per-rank array  (num_var, nz, ny_slab, nx, n^3)  [Euler]  /  (num_var, ny_slab, nx, n^2)  [SW]
 * (wx_factory/geometry/cubed_sphere_3d.py:186-190, cubed_sphere_2d.py:161-170).
 */
#ifndef WXB200_H
#define WXB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WXB_VERSION 1
#define WXB_MAX_SLABS 8
#define WXB_MAX_N 6

#define WXB_OK 0
#define WXB_EINVAL (-1)  /* bad argument / unsupported size          */
#define WXB_ECUDA (-2)   /* a CUDA runtime call or kernel failed     */
#define WXB_ENOMEM (-3)  /* device or host allocation failed         */
#define WXB_ESTATE (-4)  /* call sequence error (e.g. finish before traces) */

#define WXB_EQ_EULER3D 0
#define WXB_EQ_SHALLOW_WATER 1

#define WXB_JVP_FD 0  /* dt * (RHS(Q + eps v) - rhs0) / eps                  */
#define WXB_JVP_RAT 1 /* v - 0.5 * dt * (RHS(Q + eps v) - rhs0) / eps        */

typedef struct wxb_handle wxb_handle;

typedef struct {
    int32_t equations;             /* WXB_EQ_*                                                        */
    int32_t n;                     /* num_solpts, 2..WXB_MAX_N                                        */
    int32_t nz;                    /* elements per column (1 for shallow water)                       */
    int32_t nx;                    /* elements per slab row (= elements per panel side)               */
    int32_t nslabs;                /* slabs owned by this handle                                      */
    int32_t slab_ny[WXB_MAX_SLABS];/* rows of elements in each slab                                   */
    int32_t nghost;                /* ghost (column,face) trace records received from other handles   */
    int32_t nsend;                 /* (column,face) trace records this handle packs for other handles */
    int32_t nxform;                /* rows of the panel-edge vector transform table                   */
    int32_t advection_only;        /* Euler case_number < 13: RHS is identically zero                 */
    int32_t has_damping;           /* Euler: Rayleigh damping arrays present (DCMIP 2-1 / 2-2)        */
    int32_t has_topo;              /* SW: topography arrays present (Williamson 5)                    */
    double damp_tau0;
} wxb_config;

/* All pointers are HOST pointers; wxb_create copies what it needs to the device.
 * E = local elements (sum over slabs of nz*ny_s*nx), C = local columns (sum of ny_s*nx),
 * P = n^3 (n^2 for SW), F = n^2 (n for SW). */
typedef struct {
    /* 1-D operator blocks (DFROperators, wx_factory/geometry/operators.py:55-80,144-145) */
    const double* ext_w;  /* [n]    */
    const double* ext_e;  /* [n]    */
    const double* diff;   /* [n*n]  diff_solpt, row-major: out[i'] = sum_i diff[i'*n+i] f[i] */
    const double* corr;   /* [n*2]  correction, column 0 = min-side face, 1 = max-side       */
    const double* hfilt;  /* [n*n]  highfilter (Euler only)                                  */
    /* interior metric, local element order */
    const double* sqrtg;      /* [E*P]                                   */
    const double* h;          /* Euler [6][E*P] (11,12,13,22,23,33); SW [3][E*P] (11,12,22) */
    const double* chris;      /* Euler [27][E*P]; SW [8][E*P]            */
    const double* inv_dzdeta; /* Euler [E*P]                             */
    const double* damp;       /* Euler, optional [4][E*P]: profile, u1ref, u2ref, u3ref */
    const double* topo;       /* SW, optional [3][E*P]: hsurf, dzdx1, dzdx2             */
    /* interface metric once per physical face, slab after slab:
     * Euler comps = 4 (sqrtG, h^{d1}, h^{d2}, h^{d3}); SW comps = 3 (sqrtG, H^{1d}.. see problem.py)
     *   fm_i: per slab [comps][nz][ny_s][nx+1][F];  fm_j: [comps][nz][ny_s+1][nx][F];  fm_k: [comps][nz+1][ny_s][nx][F] */
    const double* fm_i;
    const double* fm_j;
    const double* fm_k;   /* Euler only */
    const double* topo_itf; /* SW optional: hsurf on faces: i-faces then j-faces, two-sided values
                               [ (own side, neighbour side) ... ] see csrc/sw2d.cu */
    /* horizontal connectivity, pull model: for face f in (W,E,S,N) of local column c the neighbour trace of
     * trace field v (Euler: the 5 variables + log(rho theta); SW: the 3 variables), level k, face point pt lives at
 *   trace[((base + v*vstride + k*kstride) * F) + pt']:
     *   nbr[(f*C + c)*4 + {0,1,2,3}] = base, kstride, vstride, flags
     *   flags bit0: reverse the along-edge node index; bits 8..: 1 + row of `xform` (0 = identity) */
    const int32_t* nbr;     /* [4*C*4] */
    const double* xform;    /* [nxform][n][4]: 2x2 matrix (m11,m12,m21,m22) applied to (var1,var2) per along-edge node */
    const int32_t* send;    /* [nsend]: col*4 + face of each record to pack, in send-buffer order */
} wxb_static_arrays;

int wxb_version(void);

int wxb_create(const wxb_config* cfg, const wxb_static_arrays* arrays, int device, wxb_handle** out);
int wxb_destroy(wxb_handle* h);
const char* wxb_last_error(const wxb_handle* h); /* h may be NULL: last create error */

/* 1 if the handle runs the fused kernel (Euler: euler_sweep.cu, shallow water: sw_tile.cu), 0 if the traces + main pair.
 * Chosen at create: env WXB_EULER_PATH / WXB_SW_PATH = sweep | tile | split, else the path that measured faster for the
 * order (no reference counterpart: the reference has one numpy path). */
int wxb_handle_fused(const wxb_handle* h);

/* sizes, in doubles */
int64_t wxb_state_size(const wxb_handle* h);      /* num_var * E * P                                  */
int64_t wxb_send_size(const wxb_handle* h);       /* nsend  * ntr * nz * F  (ntr trace fields: Euler 6, SW 3) */
int64_t wxb_recv_size(const wxb_handle* h);       /* nghost * ntr * nz * F                             */
double* wxb_send_buffer(wxb_handle* h);           /* device pointer, contiguous, record-major          */
double* wxb_recv_buffer(wxb_handle* h);           /* device pointer (the ghost tail of the trace array) */

/* One whole evaluation on a handle without remote neighbours (nghost == 0): dq = RHS(q). */
int wxb_rhs(wxb_handle* h, const double* q, double* dq, void* stream);

/* out = JVP(v) around q with cached rhs0 = RHS(q) (mode WXB_JVP_*), base+perturbed fused in one pass:
 * the perturbed state q + eps*v is formed in registers, the epilogue applies dt*(.-rhs0)/eps [and v - 0.5*..]. */
int wxb_jvp(wxb_handle* h, const double* q, const double* v, const double* rhs0, double dt, double eps, int mode,
            double* out, void* stream);

/* Split form for handles with remote neighbours.  wxb_traces extrapolates q (+ eps*v if v != NULL) to all
 * faces and packs the records listed in `send` into wxb_send_buffer; the caller moves them into the peers'
 * wxb_recv_buffer (NCCL / peer copy), then calls wxb_finish with the same q, v, eps.
 * part: 0 = all elements, 1 = only elements that do not read ghost records (may run before the exchange
 * has completed), 2 = only elements that do. */
int wxb_traces(wxb_handle* h, const double* q, const double* v, double eps, void* stream);
int wxb_finish(wxb_handle* h, const double* q, const double* v, const double* rhs0, double dt, double eps,
               int mode /* -1 = plain RHS */, int part, double* out, void* stream);

/* Halo exchange inside the library (multi-GPU: one handle per process / GPU).  Replaces the MPI neighbour all-to-alls of
 * ProcessTopology.start_exchange_scalars / start_exchange_vectors (wx_factory/process_topology.py:269-386) and the
 * start_communication / end_communication pair of the RHS (wx_factory/rhs/rhs_dfr.py:141-201).
 *   wxb_comm_unique_id : rank 0 fills a WXB_NCCL_ID_BYTES buffer (ncclGetUniqueId) and ships it to the other ranks by any
 *                        means (MPI_Bcast, torch.distributed, a file)
 *   wxb_comm_init      : every rank joins with (id, rank, nranks) and says which contiguous runs of its send records go
 *                        to which rank, and which runs of its ghost records come from which rank (record = ntr*nz*F
 *                        doubles; `first` / `count` in records, runs in buffer order)
 *   wxb_exchange       : ncclGroupStart; ncclSend / ncclRecv per peer; ncclGroupEnd on `stream` (send buffer -> the
 *                        peers' ghost records)
 *   wxb_eval_dist      : one whole evaluation on such a handle: boundary traces, the exchange on the handle's own
 *                        communication stream overlapped with the part that reads no ghost record, then the rest.
 *                        mode -1: out = RHS(q) (v, rhs0 ignored); WXB_JVP_*: as wxb_jvp.
 * NCCL is loaded with dlopen("libnccl.so.2") at the first call; the library has no link-time dependency on it. */
#define WXB_NCCL_ID_BYTES 128
int wxb_comm_unique_id(void* id_out);
int wxb_comm_init(wxb_handle* h, const void* id, int rank, int nranks, int n_send_peers, const int32_t* send_peer,
                  const int32_t* send_first, const int32_t* send_count, int n_recv_peers, const int32_t* recv_peer,
                  const int32_t* recv_first, const int32_t* recv_count);
int wxb_exchange(wxb_handle* h, void* stream);
int wxb_eval_dist(wxb_handle* h, const double* q, const double* v, const double* rhs0, double dt, double eps, int mode,
                  double* out, void* stream);

/* Same as wxb_rhs / wxb_jvp with HOST buffers: H2D copy, kernels, D2H copy, synchronise.  For the 3-D Euler
 * equations the state moves in chunks of vertical levels on three streams, so that the upload of one chunk, the
 * kernels of the previous one and the download of the one before overlap (env WXB_HOST_CHUNKS, default 24;
 * page-locked host buffers are needed for the copies to run asynchronously).
 * On a handle WITH remote neighbours these calls need wxb_comm_init first (WXB_ESTATE otherwise) and are collective
 * like wxb_eval_dist: on the fused Euler path with up to two ranks the records other ranks need are formed and
 * exchanged chunk by chunk behind the uploads (the exchange of a chunk, the sweep of the previous one and the
 * download of the one before overlap); otherwise the whole arrays go up, wxb_eval_dist runs, the whole result comes
 * down (env WXB_HOST_DIST=pipe|whole overrides the choice; with more ranks copying at once the measured boxes are
 * bound by host bandwidth and the overlap buys nothing). */
int wxb_rhs_host(wxb_handle* h, const double* q_host, double* dq_host);
int wxb_jvp_host(wxb_handle* h, const double* q_host, const double* v_host, const double* rhs0_host, double dt,
                 double eps, int mode, double* out_host);

/* Linearisation point kept on the device: what MatvecOp / MatvecOpRat capture at construction
 * (wx_factory/solvers/matvec.py:7-34,71-73; built once per time step by wx_factory/integrators/ros2.py:29 and the
 * EPI integrators, then applied once per Krylov iteration).
 *   wxb_linearize_host : rhs_host = RHS(q_host), and (q, RHS(q)) stay on the device as the linearisation point
 *   wxb_bind_host      : upload a (q, rhs0) pair computed elsewhere as the linearisation point
 *   wxb_jvp_bound_host : the matvec at that point; only v goes up and the product comes down
 * The point stays valid until the next wxb_linearize_host / wxb_bind_host on the handle (wxb_rhs_host and
 * wxb_jvp_host do not disturb it).  WXB_ESTATE without one. */
int wxb_linearize_host(wxb_handle* h, const double* q_host, double* rhs_host);
int wxb_bind_host(wxb_handle* h, const double* q_host, const double* rhs0_host);
int wxb_jvp_bound_host(wxb_handle* h, const double* v_host, double dt, double eps, int mode, double* out_host);

/* Device time of the last evaluation's kernels in ms: [0] traces(+pack), [1] main, [2] total. */
int wxb_timings(wxb_handle* h, float out_ms[3]);

/* The same three durations for the most recent evaluations (at most max_evals, at most 256 are kept), oldest
 * first: out_ms[3*i + {0,1,2}], out_mode[i] = -1 (RHS) / WXB_JVP_* (nullable).  Returns the number of rows
 * written, or a negative WXB_E* code.  Synchronises with the last evaluation.  Replaces the reference's
 * per-call timestamp lists (wx_factory/rhs/rhs.py:78-79,190-213). */
int wxb_timing_history(wxb_handle* h, int max_evals, float* out_ms, int* out_mode);

/* ---------------------------------------------------------------------------------------------------------------
 * Per-step operators around the right-hand side (csrc/stepops.cu): the callers' side of the boundary.
 * ------------------------------------------------------------------------------------------------------------- */

/* Per-step exponential modal filter of the 3-D state:  out = ((sqrtG q) @ expfilter_new) * inv_sqrtG
 * <- DFROperators.apply_filter_3d (wx_factory/geometry/operators.py:257-261, applied once per step by
 * wx_factory/simulation/simulation.py:148 when expfilter_apply = 1).  `filt` is the (n x n) nodal filter
 * DFROperators.expfilter (operators.py:110-112, make_filter :215-237), row-major; the three 1-D passes replace the
 * dense (n^3 x n^3) product.  out may alias q. */
int wxb_set_expfilter(wxb_handle* h, const double* filt);
int wxb_expfilter(wxb_handle* h, const double* q, double* out, void* stream);

/* Vector algebra of the integrators and Krylov solvers on plain device arrays (no handle).  All asynchronous on
 * `stream`; coefficient arrays are HOST pointers read before the call returns. */
#define WXB_VEC_MAX_ROWS 64

/* out = ((a x) + (b y)) + ((c z) dt), each operation rounded on its own: the bit pattern of numpy's
 * a*x + b*y + c*z*dt  <- Tvdrk3.__step__ (wx_factory/integrators/tvdrk3.py:13-17), Ros2.__prestep__'s
 * b = A(Q) + rhs*dt (wx_factory/integrators/ros2.py:30).  y may be NULL (term absent); a == 1 leaves x unscaled. */
int wxb_vec_lincomb3(int64_t n, double a, const double* x, double b, const double* y, double c, const double* z, double dt,
                     double* out, void* stream);

/* *count_dev = number of non-finite entries  <- the NaN check of wx_factory/simulation/simulation.py:399-408 */
int wxb_vec_nonfinite(int64_t n, const double* x, uint64_t* count_dev, void* stream);

/* out_dev[i*k + c] = sum_x V[i*ldv + x] * W[c*ldw + x], i < m <= WXB_VEC_MAX_ROWS, k in {1, 2}
 * <- V[ilow:j,:n] @ V[j,:n] and the norms of KIOPS (wx_factory/solvers/kiops.py:190-205), PMEX
 * (wx_factory/solvers/pmex.py:150-230), Q[:j] @ Q[j-2:j].T of the 1-sync FGMRES (wx_factory/solvers/fgmres.py:38).
 * Two-stage fixed-order reduction (bit-reproducible run to run); scratch_dev holds wxb_vec_multidot_scratch(m, k)
 * doubles.  The result stays on the device: the caller all-reduces it over ranks (the reference's comm.allreduce,
 * fgmres.py:41 / kiops.py:193) and reads it back -- the one synchronisation of the step. */
int64_t wxb_vec_multidot_scratch(int m, int k);
int wxb_vec_multidot(int64_t n, int m, const double* V, int64_t ldv, int k, const double* W, int64_t ldw, double* out_dev,
                     double* scratch_dev, void* stream);

/* wxb_vec_multidot for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same inputs,
 * outputs and arithmetic as wxb_vec_multidot, but the scratch of the two-stage reduction lives in a file-scope device
 * array inside the library (592 KiB on every GPU that loads it), so the caller passes none and the call performs no
 * cudaMalloc / cudaFree / allocating library call.  Limit: m <= WXB_VEC_STATIC_MAX_ROWS (WXB_EINVAL beyond it; use
 * wxb_vec_multidot with caller scratch there).  All calls on a device share the one array: calls that overlap on the GPU
 * (different streams, no ordering between them) RACE -- order them.  Nothing needs clearing between calls: every partial
 * sum the second stage reads is rewritten by the first stage of the same call. */
#define WXB_VEC_STATIC_MAX_ROWS 64
int wxb_vec_multidot_static(int64_t n, int m, const double* V, int64_t ldv, int k, const double* W, int64_t ldw, double* out_dev,
                            void* stream);

/* Scalar reductions (fixed order; scratch_dev as for wxb_vec_multidot(1, 1)):
 *   kind 0: *out_dev = sum_i |x[i]|                  <- 1-norms of the phi vectors, kiops.py:111-116, pmex.py:91-98
 *   kind 1: *out_dev = max_{x[i] != 0} |y[i] / x[i]|, -1 if x == 0   <- stagnation test update/x of fgmres.py:257-262 */
int wxb_vec_reduce(int64_t n, int kind, const double* x, const double* y, double* out_dev, double* scratch_dev, void* stream);

/* out = (b0 y + sum_{i<m} coef[i] V[i]) * s   <- V[j] - V[ilow:j].T @ H (kiops.py:199), V[j] / nrm (:208),
 * beta * F[:j,0] @ V[:j,:n] (:278,286), x + Z.T @ y of FGMRES (fgmres.py:246-249).  y may be NULL; out may alias y. */
int wxb_vec_combine(int64_t n, int m, const double* coef_host, const double* V, int64_t ldv, double b0, const double* y, double s,
                    double* out, void* stream);

/* The row updates of _ortho_1_sync_igs (wx_factory/solvers/fgmres.py:60-69) in one pass over the first m = j-2 rows:
 * q2 -= Q[:m].T s2 ; q1 -= Q[:m].T s1 ; q2 /= norm ; q1 -= q2 r ; q1 /= norm   (q2 = Q[j-2], q1 = Q[j-1]). */
int wxb_vec_gs1sync(int64_t n, int m, const double* s2_host, const double* s1_host, const double* Q, int64_t ldq, double norm,
                    double r, double* q2, double* q1, void* stream);

#ifdef __cplusplus
}
#endif
#endif
