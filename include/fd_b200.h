/*
 * fd_b200.h -- C ABI of libfd_b200.so: the B200 (sm_100a) kernels behind the
 * time-stepping hot path of the `finite_difference` Python library.
 *
 * The reference (Ephraim-Bryski/Finite-Difference 0.0.3) has no FFI: its hot
 * path is numpy code inside src/finite_difference/FD.py.  Each entry point
 * below therefore cites the FD.py lines whose arithmetic it replaces; the
 * Python package `finite_difference` (finite-difference_b200/) binds them
 * with ctypes.
 *
 * Conventions
 *   - plain pointers and sizes only; all data pointers are DEVICE pointers
 *     owned by the caller unless stated otherwise; descriptor structs are
 *     HOST pointers, read during the call only;
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream);
 *     every call is asynchronous on that stream;
 *   - every function returns 0 on success, non-zero on error, and
 *     fd_last_error() then describes the failure (thread-local string);
 *   - dtype suffix: _f64 = IEEE double (the reference's only type),
 *     _f32 = IEEE float (additive option);
 *   - arithmetic mirrors numpy's order of operations exactly: products and
 *     sums are separately rounded (no FMA contraction), stencil sums start
 *     from 0.0 and accumulate in the user's point order;
 *   - "unknown" is NaN, as in the reference (FD.py:310).
 *
 * Device-side conditions are reported through caller-owned device words:
 *   flags[FD_FLAG_INF]      != 0  a der() produced +-inf        (FD.py:1112-1113)
 *   flags[FD_FLAG_OVERRIDE] != 0  a known value hit a known one (FD.py:565-572)
 * and NaN counters (one int64 per stored time level, FD.py:528-538).
 */
#ifndef FD_B200_H
#define FD_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FD_ABI_VERSION 6

#define FD_MAX_POINTS 16 /* sample points per stencil                     */
#define FD_MAX_TERMS 4   /* der() terms summed by one fused statement     */
#define FD_MAX_POST 4    /* scalar/coefficient operations after the sum   */

#define FD_FLAG_INF 0
#define FD_FLAG_OVERRIDE 1
#define FD_FLAG_HALO_TIMEOUT 2 /* fd_halo_wait gave up (multi-GPU only) */
#define FD_NUM_FLAGS 4

/* One Stencil applied along one axis: integer index shifts (FD.py:1028-1044),
 * the 2-decimal weights solved on the host (FD.py:938-951) and the divisor
 * dx**der_order (FD.py:1107-1110).  Passed by value to the kernels, i.e. it
 * lives in the constant bank for the duration of the launch. */
typedef struct fd_stencil {
  int32_t npts;
  int32_t shifts[FD_MAX_POINTS];
  double weights[FD_MAX_POINTS];
  double scale;
  int32_t exact_div; /* 1: bit-exact quotient (the library picks the cheapest
                        exact method); 0: multiply by 1/scale (<= 1 ulp off);
                        2: force the IEEE division instruction sequence       */
} fd_stencil_t;

/* element-wise operations of FieldInstant arithmetic (FD.py:794-869) */
enum fd_ewise_op {
  FD_OP_NEG = 0,   /* dst = -a                                  FD.py:836 */
  FD_OP_ADD_FF,    /* dst = a + b                               FD.py:839 */
  FD_OP_SUB_FF,    /* dst = a - b                               FD.py:842 */
  FD_OP_MUL_FF,    /* dst = a * b                               FD.py:845 */
  FD_OP_DIV_FF,    /* dst = a / b                               FD.py:848 */
  FD_OP_POW_FF,    /* dst = pow(a, b)                           FD.py:852 */
  FD_OP_ADD_FS,    /* dst = a + s                               FD.py:839,856 */
  FD_OP_SUB_FS,    /* dst = a - s  (also `s - a`, FD.py:859-860 quirk)    */
  FD_OP_MUL_FS,    /* dst = a * s                               FD.py:845,862 */
  FD_OP_DIV_FS,    /* dst = a / s                               FD.py:848 */
  FD_OP_POW_FS,    /* dst = pow(a, s)                           FD.py:852 */
  FD_OP_SQUARE,    /* dst = a * a   (numpy's a**2 fast path)              */
  FD_OP_RECIP,     /* dst = 1 / a   (numpy's a**-1 fast path)             */
  FD_OP_SQRT,      /* dst = sqrt(a) (numpy's a**0.5 fast path)            */
  FD_OP_COPY,      /* dst = a       (numpy's a**1 fast path)              */
  FD_OP_EULER      /* dst = a + b*s (two roundings)             FD.py:515-516 */
};

/* ---- library ------------------------------------------------------------ */

/* ABI version of the loaded library (== FD_ABI_VERSION). No reference
 * counterpart. */
int fd_abi_version(void);

/* Human-readable description of the last failure on this thread. */
const char* fd_last_error(void);

/* Number of kernels this library has launched since load (all threads);
 * bench.py reports it as gpu_launches. */
long long fd_launch_count(void);

/* Which kernel the last fd_fused_* call of this thread launched and with what
 * geometry, e.g. "star_lean f64 R=1 axes=3 vy=2 vz=1 chunk=8 grid=8x64x64".
 * Diagnostic: the parity tests at the benchmark's sizes assert that they ran the
 * launch configuration the benchmark times.  No reference counterpart. */
const char* fd_last_kernel(void);

/* ---- unfused operators (any expression DAG, any number of dimensions) ---- */

/* Stencil.der on one time slice viewed as (outer, n_src, inner) ->
 * (outer, n_dst, inner), the derivative axis in the middle.
 * periodic != 0: FD.py:1046-1055 (index wrap instead of np.roll copies),
 *   n_dst must equal n_src.
 * periodic == 0: FD.py:1058-1105; outputs with lo <= j < hi are computed from
 *   src[j + shift_i], the others are NaN; n_dst = n_src + {-1, 0, +1}.
 * Then / dx**order (FD.py:1107-1110) and the infinity flag (FD.py:1112). */
int fd_der_f64(void* dst, const void* src, int64_t outer, int64_t n_src, int64_t n_dst,
               int64_t inner, const fd_stencil_t* st, int periodic, int64_t lo, int64_t hi,
               void* flags, void* stream);
int fd_der_f32(void* dst, const void* src, int64_t outer, int64_t n_src, int64_t n_dst,
               int64_t inner, const fd_stencil_t* st, int periodic, int64_t lo, int64_t hi,
               void* flags, void* stream);

/* FieldInstant arithmetic on n elements (FD.py:794-869); `b` is ignored by
 * the field-scalar and unary forms, `s` by the field-field forms. dst may
 * alias a or b. */
int fd_ewise_f64(int op, void* dst, const void* a, const void* b, double s, int64_t n, void* stream);
int fd_ewise_f32(int op, void* dst, const void* a, const void* b, double s, int64_t n, void* stream);

/* Field.__set_data with allow_override=False (FD.py:541-576): merge the
 * freshly computed slice `fresh` into the stored level `stored` (already
 * holding NaN or boundary values): known fresh values are written (raising
 * FD_FLAG_OVERRIDE if the stored one was known too), NaN fresh values keep
 * what is stored.  `nan_count` (device int64*, may be NULL) is incremented by
 * the number of NaNs left in `stored` (FD.py:528-529). */
int fd_merge_store_f64(void* stored, const void* fresh, int64_t n, void* nan_count, void* flags, void* stream);
int fd_merge_store_f32(void* stored, const void* fresh, int64_t n, void* nan_count, void* flags, void* stream);

/* Fill n elements with `value` (NaN to make a level "unknown", FD.py:310). */
int fd_fill_f64(void* dst, double value, int64_t n, void* stream);
int fd_fill_f32(void* dst, double value, int64_t n, void* stream);

/* Write one boundary face into a stored level viewed as (outer, n, inner):
 * stored[o, index, i] = face[o, i] wherever face is known (set_BC with
 * allow_override=True, FD.py:361-385 and 541-576). */
int fd_face_store_f64(void* stored, const void* face, int64_t outer, int64_t n, int64_t inner, int64_t index, void* stream);
int fd_face_store_f32(void* stored, const void* face, int64_t outer, int64_t n, int64_t inner, int64_t index, void* stream);

/* Count NaNs of n elements into *nan_count (device int64*, incremented). */
int fd_count_nan_f64(const void* src, int64_t n, void* nan_count, void* stream);
int fd_count_nan_f32(const void* src, int64_t n, void* nan_count, void* stream);

/* ---- fused statement: der terms + scalar arithmetic + Euler chain + BCs ---- */

/* One der() term of the sum. The source slice has the target's dimensions
 * except along `axis`, where it has n_src entries (staggered fixed axes differ
 * by one). */
typedef struct fd_term {
  const void* src;
  int32_t axis;      /* 0,1,2 in the collapsed (n0,n1,n2) view, n2 contiguous
                        (fewer than three spatial axes: leading extents are 1) */
  int32_t periodic;
  int32_t lo, hi;    /* computed output range along axis when not periodic    */
  int32_t n_src;
  int32_t subtract;  /* 0: sum += term, 1: sum -= term (first term: ignored)   */
  int32_t has_pre;   /* multiply the term by `pre` before summing (k*der(..))  */
  int32_t negate;    /* unary minus on the term before summing                 */
  /* More than three spatial axes (FD.py:1007-1012 applies der to any axis of an
   * N-D array): dimension 0 of the collapsed view is then the row-major merge of
   * all leading axes, and a term along one of THEM sets axis = 0, fold_mod = that
   * axis' extent and fold_div = the product of the extents of the merged axes
   * behind it; its index along the axis is (i0 / fold_div) % fold_mod.  The
   * source must have the target's extents (n_src == fold_mod) and no boundary
   * face may lie on dimension 0.  fold_mod == 0: dimension 0 is one axis. */
  int32_t fold_div, fold_mod;
  double pre;
  fd_stencil_t st;
} fd_term_t;

enum fd_post_op {
  FD_POST_MUL = 0,  /* v = v * s        */
  FD_POST_ADD,      /* v = v + s        */
  FD_POST_SUB,      /* v = v - s        */
  FD_POST_DIV,      /* v = v / s        */
  FD_POST_NEG,      /* v = -v           */
  FD_POST_MUL_FIELD /* v = v * coef[i]  (ConstantField.always, FD.py:759-762;
                       the product is commutative so coef*v is identical) */
};

typedef struct fd_post {
  int32_t op;
  int32_t _pad;
  double s;
  const void* coef; /* FD_POST_MUL_FIELD: slice with the target's dimensions */
} fd_post_t;

/* `Field.dot...dot.assign_update(post(sum of terms))` followed by `depth`
 * time_integrate_update calls up the .dot chain, executed in ONE pass:
 *   rhs       = post(sum_k term_k)                   FD.py:990-1122, 794-869
 *   dot_out   = rhs                     (optional)   FD.py:478-498
 *   out[0]    = merge(prev[0] + rhs*dt,    faces[0]) FD.py:500-538, 541-576
 *   out[1]    = merge(prev[1] + out[0]*dt, faces[1]) (depth == 2)
 * prev[0] == NULL is the assign_update alone (an auxiliary field such as a flux,
 * or a .dot that is not integrated in this step; FD.py:478-498 with the store of
 * FD.py:541-576):  out[0] = merge(rhs, faces[0]),  depth = 1, dot_out unused.
 * Boundary faces hold the values the reference pre-populates at time t+1 with
 * set_BC (FD.py:361-385); face[l][d][side] is an array over the two other
 * dimensions in order, or NULL.  Periodic axes wrap the index in-kernel,
 * fixed axes predicate on [lo, hi).
 * Multi-GPU: a slab carries ghost planes along its decomposed dimension; the
 * caller exchanges them (fd_halo_push below, or send/recv), describes that dimension as
 * non-periodic with [lo, hi) = the planes it owns and restricts execution
 * with box_lo / box_hi. */
typedef struct fd_fused {
  int32_t n[3];
  int32_t nterms;
  fd_term_t term[FD_MAX_TERMS];
  int32_t npost;
  int32_t depth;      /* 1 or 2 stored levels */
  fd_post_t post[FD_MAX_POST];
  double dt;
  void* dot_out;
  const void* prev[2];
  void* out[2];
  const void* face[2][3][2];
  void* nan_count[2]; /* device int64*, incremented by NaNs left in out[l]   */
  void* flags;        /* device int64[FD_NUM_FLAGS]                          */
  /* Sub-box execution (slab decomposition, interior/rim split): only points
   * with box_lo[d] <= index[d] < box_hi[d] for every d are computed and
   * stored; an all-zero box means the whole slice.  With ghost planes around
   * a slab the domain faces do not sit at index 0 / n-1: rim_lo[d] / rim_hi[d]
   * give the index along d where face[l][d][0] / face[l][d][1] apply. */
  int32_t box_lo[3], box_hi[3];
  int32_t rim_lo[3], rim_hi[3];
} fd_fused_t;

int fd_fused_step_f64(const fd_fused_t* desc, void* stream);
int fd_fused_step_f32(const fd_fused_t* desc, void* stream);

/* ---- fused staggered pair: the velocity updates and the pressure update of a
 *      cell/edge leapfrog step in ONE pass ---------------------------------- */

/* For every listed axis a (user script):
 *     v_a.dot.assign_update(c_up * up.der(p.prev, a));  v_a.time_integrate_update()
 * and then
 *     p.dot.assign_update(c_down * (sum_a down.der(v_a.new, a)));  p.time_integrate_update()
 * i.e. 2*(naxes+1) statements of FD.py:478-538 with the der calls of
 * FD.py:990-1122 between them.  The kernel recomputes v_a.new at the second
 * sample point instead of reading it back, so each field is read once and
 * written once.  `up` is a cell->edge stencil with shifts (-1, 0), `down` an
 * edge->cell stencil with shifts (0, +1); every axis must be periodic or be the
 * decomposed axis of a slab (ghost planes, box_lo/box_hi = owned planes). */
typedef struct fd_pair_axis {
  const void* v_prev;
  void* v_out;
  void* vdot_out;     /* optional: v_a.dot at t+1 */
  void* v_nan_count;  /* device int64*, may be NULL */
  double c_up;
  int32_t axis;       /* 0,1,2 in the collapsed view */
  int32_t periodic;   /* 0: the axis carries ghost planes (slab) */
  int32_t subtract;   /* 1: this term is subtracted in the sum (first term: ignored) */
  int32_t _pad;
  fd_stencil_t up, down;
} fd_pair_axis_t;

typedef struct fd_pair {
  int32_t n[3];
  int32_t naxes;
  fd_pair_axis_t ax[3];
  const void* p_prev;
  void* p_out;
  void* pdot_out;     /* optional */
  void* p_nan_count;
  void* flags;
  double c_down, dt;
  int32_t box_lo[3], box_hi[3];
} fd_pair_t;

int fd_fused_pair_f64(const fd_pair_t* desc, void* stream);
int fd_fused_pair_f32(const fd_pair_t* desc, void* stream);

/* ---- fused statement with a general point-wise expression ----------------- */

/* Any FieldInstant expression (FD.py:794-869) over der terms (FD.py:990-1122),
 * stored field levels and numbers -- products of fields, powers, quotients --
 * followed by the same assign_update / time_integrate_update chain as
 * fd_fused_step, in one pass.  The expression is a postfix program evaluated
 * per point on a small value stack; every operation is the same IEEE operation
 * the unfused fd_ewise kernels perform, in the same order. */
#define FD_EXPR_MAX_FIELDS 4
#define FD_EXPR_MAX_CODE 32
#define FD_EXPR_MAX_STACK 8

enum fd_expr_op {
  FD_X_PUSH_DER = 0, /* arg: index into base.term                     */
  FD_X_PUSH_FIELD,   /* arg: index into field[]                       */
  FD_X_PUSH_IMM,     /* imm                                           */
  FD_X_NEG,          /* top = -top                                    */
  FD_X_ADD, FD_X_SUB, FD_X_MUL, FD_X_DIV, FD_X_POW,             /* b = pop, a = pop, push a op b */
  FD_X_ADD_IMM, FD_X_SUB_IMM, FD_X_MUL_IMM, FD_X_DIV_IMM, FD_X_POW_IMM, /* top = top op imm */
  FD_X_SQUARE, FD_X_RECIP, FD_X_SQRT                             /* numpy's scalar-power fast paths */
};

typedef struct fd_expr {
  fd_fused_t base;   /* extents, der terms (subtract/pre/negate and post are not
                        used: nterms may be 0, npost must be 0), Euler chain,
                        faces, counters, box */
  int32_t nfields;
  int32_t ncode;
  const void* field[FD_EXPR_MAX_FIELDS]; /* levels with the target's dimensions */
  int32_t op[FD_EXPR_MAX_CODE];
  int32_t arg[FD_EXPR_MAX_CODE];
  double imm[FD_EXPR_MAX_CODE];
} fd_expr_t;

int fd_fused_expr_f64(const fd_expr_t* desc, void* stream);
int fd_fused_expr_f32(const fd_expr_t* desc, void* stream);

/* ---- a derivative of a computed expression -------------------------------- */

/* Stencil.der accepts any FieldInstant (FD.py:990-1003), e.g.
 *     T.dot.assign_update(d.der(kappa.always * d.der(T.prev, "x"), "x"))
 *     u.dot.assign_update(-0.5 * d.der(u.prev * u.prev, "x"))
 * In the reference the inner expression is an array of its own.  Here it is
 * evaluated ONCE into a scratch level by a plain-assignment launch (`inner`:
 * depth 1, prev[0] == NULL, out[0] == NULL = "the scratch level"), and the
 * statement proper (`outer`) then reads the scratch level wherever it names a
 * NULL source (term[k].src, and field[k] of an fd_expr_t): two launches instead
 * of one per reference array operation, the same values bit for bit.
 * inner_is_expr / outer_is_expr: the descriptor is an fd_expr_t (1) or an
 * fd_fused_t (0).  `scratch` holds scratch_elems values of the entry point's
 * type; the inner expression must fit (n[0]*n[1]*n[2] values). */
int fd_fused_nested_f64(int inner_is_expr, const void* inner, int outer_is_expr, const void* outer, void* scratch,
                        long long scratch_elems, void* stream);
int fd_fused_nested_f32(int inner_is_expr, const void* inner, int outer_is_expr, const void* outer, void* scratch,
                        long long scratch_elems, void* stream);

/* fd_fused_nested_f64 for callers that have no device memory to give at this
 * point.  It has no counterpart in the reference (numpy allocates the inner
 * array afresh on every call) and no counterpart among the other entry points,
 * which all work on buffers the caller owns.  Meant for hosts that embed the
 * stepper and may not allocate while they step: an application that handed the
 * whole GPU to a memory pool of its own when it started, or one whose time
 * loop must stay free of cudaMalloc because that call synchronises the device
 * and can still fail late, when memory has become fragmented.  Such a caller
 * still owns every input and output level; only the scratch level comes from
 * the library: a fixed array of FD_STATIC_SCRATCH_BYTES that is part of the
 * loaded module (each GPU that loads the library carries it until the process
 * ends).  The function allocates nothing, on the host or on the device.
 *   - The inner expression may have at most FD_STATIC_SCRATCH_BYTES / 8 values
 *     (2 Mi doubles, e.g. a 1448 x 1448 or a 128^3 level); anything larger is
 *     refused with an error before any launch.
 *   - The first call on each device looks up the array's device address
 *     (cudaGetSymbolAddress, with that device current) and keeps it; make that
 *     first call outside any stream capture.
 *   - All calls on a device share the one array.  Nothing in it has to be
 *     cleared between calls -- the inner launch writes every value the outer
 *     launch reads -- but two calls whose launches can overlap on the GPU
 *     (different streams, no ordering between them) race on it: order them. */
#define FD_STATIC_SCRATCH_BYTES (16 * 1024 * 1024)
int fd_fused_nested_static_f64(int inner_is_expr, const void* inner, int outer_is_expr, const void* outer, void* stream);

/* ---- halo exchange over peer memory (multi-GPU slabs) --------------------
 *
 * No reference counterpart: the reference is a single-process numpy program.
 * These calls replace the isend/irecv pair of finite_difference/_dist.py on one
 * NVLink/NVSwitch box: neighbouring ranks (one process per GPU) write each
 * other's ghost planes directly.
 *
 * fd_ipc_export: CUDA-IPC handle of the allocation `dev_ptr` lies in (64 bytes
 *   into handle_out) and the pointer's offset inside it.  The caller ships both
 *   to the peer process (any host channel).
 * fd_ipc_open:   maps a peer's allocation into this process (peer access is
 *   enabled on demand); *base_out + offset is then a device pointer into the
 *   peer's memory.  Open an allocation once per process.
 * fd_halo_push:  copies `nbytes` (multiple of 16) from src_lo to peer_dst_lo and
 *   from src_hi to peer_dst_hi (either side may be NULL/NULL), then stores
 *   `value` into the peers' flag words with system-scope release semantics.
 * fd_halo_wait:  one polling thread on `stream` until the int64 words flag_a and
 *   flag_b (either may be NULL) are >= value; after timeout_ns it raises
 *   flags[FD_FLAG_HALO_TIMEOUT] and returns instead of hanging. */
#define FD_IPC_HANDLE_BYTES 64
int fd_ipc_export(const void* dev_ptr, void* handle_out, long long* offset_out);
int fd_ipc_open(const void* handle, void** base_out);
int fd_ipc_close(void* base);
int fd_halo_push(const void* src_lo, void* peer_dst_lo, const void* src_hi, void* peer_dst_hi, long long nbytes,
                 void* peer_flag_lo, void* peer_flag_hi, long long value, void* stream);
int fd_halo_wait(const void* flag_a, const void* flag_b, long long value, long long timeout_ns, void* flags, void* stream);

/* fd_halo_exchange: push + publish + wait of up to FD_MAX_LINKS fields in ONE
 *   launch.  Every block copies its share, the block that finishes last publishes
 *   `value` in the neighbours' flag words and then waits for this rank's own
 *   words (same time-out behaviour as fd_halo_wait).  `counter` is a device
 *   uint32 the caller zeroes once; the kernel leaves it zero.  Launches that
 *   share a counter must be ordered (same stream). */
#define FD_MAX_LINKS 4
typedef struct fd_halo_link {
  const void* src_lo;      /* this rank's lowest owned planes  -> peer_dst_lo (previous rank's high ghosts) */
  void* peer_dst_lo;
  const void* src_hi;      /* this rank's highest owned planes -> peer_dst_hi (next rank's low ghosts)      */
  void* peer_dst_hi;
  long long nbytes;        /* per side, a multiple of 16 */
  void* peer_flag_lo;      /* the neighbours' flag words that receive `value` */
  void* peer_flag_hi;
  const void* my_flag_lo;  /* this rank's flag words to wait on (>= value)    */
  const void* my_flag_hi;
  long long value;
} fd_halo_link_t;
int fd_halo_exchange(const fd_halo_link_t* links, int nlinks, void* counter, long long timeout_ns, void* flags, void* stream);

/* One statement of a slab step, issued natively (the reference has neither
 * slabs nor streams): the interior box [lo+depth, hi-depth) of `desc` along the
 * decomposed dimension on the compute stream; the two boundary boxes of `depth`
 * planes and then fd_halo_exchange of `link[]` on the communication stream;
 * the events that order the two streams.  Afterwards the compute stream has
 * waited for the boundary boxes but not for the exchange: fd_seq_wait_halo
 * makes a stream wait for the exchange issued last.  `seq` (fd_seq_create)
 * holds the events and the ticket counter; one per process and device. */
typedef struct fd_slab_step {
  int32_t kind;            /* descriptor type: 0 fd_fused_t, 1 fd_pair_t, 2 fd_expr_t */
  int32_t dim;             /* decomposed dimension of the collapsed view */
  int32_t lo, hi;          /* owned planes along it: [lo, hi), hi - lo >= 3 * depth */
  int32_t depth;           /* ghost planes on either side */
  int32_t nlinks;          /* fields whose new boundary planes travel (0: none) */
  const void* desc;        /* box_lo/box_hi of the other dimensions are taken from it */
  fd_halo_link_t link[FD_MAX_LINKS];
  long long timeout_ns;
  void* flags;
  void* compute_stream;
  void* comm_stream;
} fd_slab_step_t;
int fd_seq_create(void** seq_out);
int fd_seq_destroy(void* seq);
int fd_seq_wait_halo(void* seq, void* stream);
int fd_slab_step_f64(void* seq, const fd_slab_step_t* step);
int fd_slab_step_f32(void* seq, const fd_slab_step_t* step);

/* fd_fused_step for a list of boxes in one call (the interior box and the rim
 * slabs of a fixed-boundary statement, FD.py:1058-1105 ranges): boxes[i] =
 * {lo0, hi0, lo1, hi1, lo2, hi2}; box 0 runs on `stream`, the others beside it
 * on `side_stream` (NULL: all on `stream`), and `stream` has waited for them
 * when the call returns.  `seq` provides the two events. */
int fd_fused_step_boxes_f64(void* seq, const fd_fused_t* desc, int nboxes, const int32_t* boxes, void* stream, void* side_stream);
int fd_fused_step_boxes_f32(void* seq, const fd_fused_t* desc, int nboxes, const int32_t* boxes, void* stream, void* side_stream);

#ifdef __cplusplus
}
#endif
#endif /* FD_B200_H */
