/*
 * xlb_b200.h -- C ABI of libxlb_b200.so: hand-written sm_100a kernels for the per-timestep LBM
 * update of XLB (hsalehipour/XLB), the path BASELINE.json:north_star names.
 *
 * The reference is 100% Python (JAX / Warp JIT); it has NO FFI of its own, so every entry point
 * below cites the reference *operator* it replaces (file:line relative to the reference tree).
 * The reference-side binding a maintainer would add is the ctypes stub shown in INTEGRATION.md.
 *
 * Conventions
 *   - every function returns 0 on success, a positive cudaError_t or a negative XLB_E* code;
 *     xlb_last_error() gives the message of the last failure on the calling thread;
 *   - all data pointers are BORROWED device pointers (e.g. torch ``tensor.data_ptr()``); nothing
 *     is retained after the call returns; kernels are enqueued on ``stream`` (a cudaStream_t
 *     passed as void*; NULL = legacy default stream) and the call does not synchronise;
 *   - fields are C-contiguous [cardinality][nx][ny][nz] (z fastest), population l starts at
 *     element l * q_stride (q_stride >= nx*ny*nz lets the host pad between populations);
 *   - a 2-D lattice (D2Q9) field [q][NX][NY] is passed as nx=1, ny=NX, nz=NY;
 *   - there is no CPU fallback anywhere in this library.
 */
#ifndef XLB_B200_H
#define XLB_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define XLB_B200_ABI_VERSION 4

/* element types (xlb/precision_policy.py:13-54) */
enum { XLB_F16 = 0, XLB_F32 = 1, XLB_F64 = 2, XLB_U8 = 3 };
/* collision models (xlb/operator/stepper/nse_stepper.py:79-87) */
enum { XLB_BGK = 0, XLB_KBC = 1, XLB_SMAGORINSKY_LES_BGK = 2 /* smagorinsky_les_bgk.py:44-66 */ };
/* boundary-condition kinds (xlb/operator/boundary_condition/bc_*.py) */
enum {
    XLB_BC_NONE = 0,
    XLB_BC_FULLWAY = 1,        /* bc_fullway_bounce_back.py:50-56   COLLISION step */
    XLB_BC_EQUILIBRIUM = 2,    /* bc_equilibrium.py:72-80           STREAMING step */
    XLB_BC_HALFWAY = 3,        /* bc_halfway_bounce_back.py:116-134 STREAMING step */
    XLB_BC_DONOTHING = 4,      /* bc_do_nothing.py:50-54            STREAMING step */
    XLB_BC_ZOUHE_VELOCITY = 5, /* bc_zouhe.py:290-304, bc_type="velocity" */
    XLB_BC_ZOUHE_PRESSURE = 6, /* bc_zouhe.py:290-304, bc_type="pressure" */
    XLB_BC_EXTRAPOLATION = 7,  /* bc_extrapolation_outflow.py:106-145 (streaming apply + post-collision assemble) */
    XLB_BC_REGULARIZED_VELOCITY = 8, /* bc_regularized.py:116-135 (Zou-He + Latt regularisation), bc_type="velocity" */
    XLB_BC_REGULARIZED_PRESSURE = 9  /* bc_regularized.py:116-135, bc_type="pressure" */
};
/* error codes */
enum { XLB_EINVAL = -1, XLB_EUNSUPPORTED = -2, XLB_EALIGN = -3 };
/* kernel variants for xlb_step (0 picks the tuned default) */
enum { XLB_VARIANT_AUTO = 0, XLB_VARIANT_SCALAR = 1, XLB_VARIANT_VECTOR = 2, XLB_VARIANT_GENERIC = 3, XLB_VARIANT_BULK = 4 };

/* One row per BC id (index = the uint8 stored in bc_mask; row 0 unused).  Built on the host and
 * uploaded once; replaces the per-stepper BC specialisation of nse_stepper.py:351-380. */
typedef struct XlbBcEntry {
    int32_t kind;         /* XLB_BC_* */
    int32_t pv_dtype;     /* element type of pv (XLB_F16/F32/F64) */
    int32_t normal[3];    /* ExtrapolationOutflowBC.normal, kernel axes (bc_extrapolation_outflow.py:78-92) */
    int32_t pv_mode;      /* velocity types: 0 = pv holds the velocity VECTOR (d components, the JAX convention, bc_zouhe.py:119-120);
                             1 = pv holds ONE scalar per cell, the inflow speed along the inward normal, u = -value * normal (the
                             Warp convention of a per-index profile, bc_zouhe.py:332-337, helper_functions_bc.py:417-473) */
    int64_t pv_stride[4]; /* strides (elements) of pv over (component, x, y, z); 0 = broadcast */
    const void* pv;       /* Zou-He prescribed values, device pointer (bc_zouhe.py:119-120, 178-215) */
    double table[27];     /* EQUILIBRIUM: feq_l (bc_equilibrium.py:74-76); HALFWAY: 6 w_l c_l.u_wall (bc_halfway_bounce_back.py:123-128) */
} XlbBcEntry;

/* Geometry + precision of one local block of the lattice. */
typedef struct XlbDomain {
    int32_t d, q;                      /* lattice: (2,9) (3,19) (3,27)  (xlb/velocity_set/*.py) */
    int32_t nx, ny, nz;                /* allocated extents (kernel axes; ghost planes included) */
    int32_t x_periodic;                /* 1: pull wraps in x (single block); 0: planes x_begin-1 / x_end exist (halo) */
    int32_t compute_dtype, store_dtype;/* precision policy (xlb/precision_policy.py:57-97) */
    int64_t q_stride;                  /* elements between populations of f fields */
    int64_t mask_q_stride;             /* elements between directions of missing_mask */
} XlbDomain;

/* Optional parameters of the collision step beyond omega (xlb_step_ex / xlb_collide_ex). */
typedef struct XlbCollisionParams {
    double smagorinsky_coef; /* SmagorinskyLESBGK(smagorinsky_coef=0.17), smagorinsky_les_bgk.py:34-42 */
    int32_t has_force;       /* != 0: ForcedCollision + ExactDifference (forced_collision.py:51-57, exact_difference_force.py:52-76) */
    int32_t bc_kind_mask;    /* OR of (1 << XLB_BC_*) over the kinds in bc_table; 0 = unknown.  Lets the library pick the leanest
                                kernel instance (e.g. without the vectorised half-way bounce-back path) */
    double force[3];         /* force_vector in KERNEL axes (a 2-D force (Fx, Fy) is passed as (0, Fx, Fy)) */
    double halfway_fraction; /* hint: fraction of the block's cells tagged with a HalfwayBounceBackBC (obstacle interiors
                                included); 0 = unknown / none.  Only steers the choice between kernel variants */
    const int64_t* bc_cells; /* optional device list (see xlb_list_bc_cells) of the linear indices, inside planes [x_begin, x_end),
                                of the cells whose BC kind needs the per-cell engine (DoNothing, Zou-He, regularized, extrapolation
                                outflow).  When given, the fused kernel leaves those cells to a second, dense launch over the list
                                instead of diverging into the per-cell engine; NULL = handle them inside the fused kernel */
    int64_t n_bc_cells;      /* entries of bc_cells */
    const uint32_t* missing_bits; /* optional device array [nx*ny*nz] from xlb_pack_missing: bit l of word `cell` = missing_mask[l][cell].
                                The boundary paths that need a cell's missing directions (half-way, Zou-He, regularized,
                                extrapolation) then read ONE word per cell instead of q bytes from q planes; NULL = read the bytes.
                                Must describe the same missing_mask that is passed to the call; 16-byte aligned */
} XlbCollisionParams;

/* Which planes (and, for the 2-D slab decomposition, which rows of each plane) one launch updates: planes
 * x_begin + k * x_stride (k < n_planes), rows y_begin + k * y_stride (k < n_rows).  n_rows == 0 means every row.
 * A stride lets ONE launch update the two edge planes of a slab (x_begin = first owned, x_stride = owned - 1, n_planes = 2). */
typedef struct XlbStepRange {
    int32_t x_begin, x_stride, n_planes;
    int32_t y_begin, y_stride, n_rows;
} XlbStepRange;

/* Halo push: the peer-memory half of the multi-GPU step (xlb/distribute/distribute.py:23-44, the lax.ppermute ring).
 * While the step kernel updates slab-axis coordinate src[k] (the first owned plane for k = 0, the last owned plane for k = 1;
 * planes of a 3-D block, rows of a D2Q9 block) it stores the populations that cross that side of the slab (c = -1 for k = 0,
 * c = +1 for k = 1 along the slab axis: velocity_set.left_indices / right_indices) ALSO into the ring neighbour's f1 buffer --
 * a device pointer of another GPU mapped into this process (xlb_ipc_import), same plane layout as f1, its own q stride -- at its ghost
 * plane / row dst[k].  Compute and transfer are one kernel: the NVLink stores are posted while the rest of the slab is being
 * updated; no pack buffer, no collective call.  Ordering between ranks is the caller's (xlb_halo_signal / xlb_halo_wait). */
typedef struct XlbHaloPush {
    void* peer_f1[2];   /* [0] left neighbour's f1, [1] right neighbour's f1; NULL = no push on that side */
    int32_t src[2];     /* my plane (row) whose outgoing populations are pushed: first owned, last owned */
    int32_t dst[2];     /* the neighbour's ghost plane (row) that receives them: its right ghost, its left ghost */
    int64_t peer_q_stride[2]; /* elements between populations in THAT neighbour's buffer (slabs of unequal thickness have unequal
                                 strides); 0 = the same as dom->q_stride */
} XlbHaloPush;

const char* xlb_last_error(void);
int xlb_abi_version(void);
int xlb_bc_entry_size(void);
int xlb_domain_size(void);
int xlb_collision_params_size(void);
/* number of kernels this library has launched from the calling process (for bench gpu_launches) */
long long xlb_launch_count(void);

/* IncompressibleNavierStokesStepper.__call__  (nse_stepper.py:227-282 JAX pull; :427-476 Warp kernel)
 * f1[:, x_begin:x_end] = collide(bc(stream(f0)))  -- one fused launch.  bc_table/bc_mask/missing_mask
 * may be NULL when n_bc == 0 (fully periodic).
 * ``timestep`` is the reference's argument of the same name (nse_stepper.py:229; its JAX path never reads it).  The result does
 * not depend on it; its parity only selects the order in which the launch walks the planes (odd: last to first), so that a
 * step starts on the planes the previous step wrote last, which are the ones still in the L2. */
int xlb_step(const XlbDomain* dom, int collision, const void* f0, void* f1, const uint8_t* bc_mask,
             const uint8_t* missing_mask, const XlbBcEntry* bc_table, int n_bc, double omega, int timestep,
             int x_begin, int x_end, int variant, void* stream);

/* Same with the extra collision parameters: collision may be XLB_SMAGORINSKY_LES_BGK, and cp->has_force adds the
 * exact-difference body force after the inner collision (nse_stepper.py:83-87).  cp == NULL behaves like xlb_step. */
int xlb_step_ex(const XlbDomain* dom, int collision, const XlbCollisionParams* cp, const void* f0, void* f1, const uint8_t* bc_mask,
                const uint8_t* missing_mask, const XlbBcEntry* bc_table, int n_bc, double omega, int timestep, int x_begin, int x_end,
                int variant, void* stream);

/* The general form of the step: the planes / rows of ``range`` (NULL = the whole block), optionally with the halo push of the
 * outgoing edge populations into the ring neighbours' buffers (``halo``, may be NULL).  With cp->bc_cells the list must cover
 * planes [x_begin, x_begin + (n_planes-1)*x_stride]; entries outside the launch's planes / rows are skipped. */
int xlb_step_range(const XlbDomain* dom, int collision, const XlbCollisionParams* cp, const XlbStepRange* range, const XlbHaloPush* halo,
                   const void* f0, void* f1, const uint8_t* bc_mask, const uint8_t* missing_mask, const XlbBcEntry* bc_table, int n_bc,
                   double omega, int timestep, int variant, void* stream);

/* The same step on ONE population buffer (in-place "AA" pattern; BASELINE.json north_star: "optionally the in-place AA pattern
 * to halve memory"; no counterpart call in the reference, whose steppers are double-buffered, nse_stepper.py:227-282).
 * parity 0 ("P" step): the buffer holds the populations in the normal layout f[l][x]; the step pulls like xlb_step and writes
 * result m of cell x to slot opp(m) of cell x + c_m -- a location the same cell has just read, so the update is race free.
 * parity 1 ("L" step): reads slot opp(l) of its own cell, writes slot m of its own cell: the layout is normal again.
 * Calls must alternate 0, 1, 0, ...; after an even number of steps the buffer equals f1 of the two-buffer scheme BIT FOR BIT.
 * Boundary conditions: none, FULLWAY and EQUILIBRIUM only (cp->bc_kind_mask must say so when n_bc > 0): the kinds that need
 * f_pre or a neighbour need the second buffer.  Collisions: BGK, KBC, SmagorinskyLESBGK without a body force.  Whole block. */
int xlb_step_inplace(const XlbDomain* dom, int collision, const XlbCollisionParams* cp, void* f, const uint8_t* bc_mask,
                     const XlbBcEntry* bc_table, int n_bc, double omega, int parity, void* stream);

/* Rank-to-rank ordering of the halo push, as two one-thread kernels on the caller's stream (capturable in a CUDA graph).
 * Every rank owns a monotonic 64-bit flag per side, written by that side's neighbour, and two local counters.
 *   xlb_halo_signal: ++*sent; store *sent to the two neighbours' flags (system-scope release after a system fence) -- enqueue it
 *                    right after the edge launch: "my push number *sent has landed in your ghost planes";
 *   xlb_halo_wait  : spin until both of MY flags are >= waited[0], then ++waited[0] -- enqueue it before the next edge launch
 *                    (step 0 passes at once: the ghost planes of the initial state are filled by the setup exchange).
 *                    ``waited`` points at TWO words: waited[1] stays 0 unless a neighbour's signal did not arrive within 30 s,
 *                    in which case the kernel gives up (the GPU is never left spinning) and stores the missed epoch there.
 * The wait of step t+1 also orders the NEXT push after the neighbour's last read of the ghost planes it overwrites, because a
 * neighbour signals only after its own edge launch of step t has finished. */
int xlb_halo_signal(uint64_t* sent, uint64_t* left_neighbours_flag, uint64_t* right_neighbours_flag, void* stream);
int xlb_halo_wait(uint64_t* waited, const uint64_t* flag_from_left, const uint64_t* flag_from_right, void* stream);

/* Peer mapping of device memory between the per-GPU processes of one node (CUDA IPC; the plumbing under XlbHaloPush).
 * export: ``handle`` receives the 64-byte cudaIpcMemHandle_t of the allocation that contains ``dptr`` and ``*offset`` the byte
 * offset of ``dptr`` inside it.  import: maps the allocation into this process (peer access enabled lazily) and returns
 * base + offset in ``*dptr`` and the mapping's base in ``*base`` (one import per distinct handle; close with xlb_ipc_close). */
int xlb_ipc_export(const void* dptr, void* handle, int64_t* offset);
int xlb_ipc_import(const void* handle, int64_t offset, void** dptr, void** base);
int xlb_ipc_close(void* base);

/* Strided batch copy on ``stream``: ``n_runs`` runs of ``run_bytes`` contiguous bytes, run r from src + r * src_stride_bytes to
 * dst + r * dst_stride_bytes, in any direction (host memory should be pinned).  One call moves an x-chunk of all q populations
 * between a host lattice and a device staging block (HostStreamedStepper) instead of q separate copies. */
int xlb_copy_runs(void* dst, int64_t dst_stride_bytes, const void* src, int64_t src_stride_bytes, int64_t run_bytes, int n_runs, void* stream);

/* Compacts the cells of planes [x_begin, x_end) whose bc id maps to a kind in ``kind_mask`` (OR of 1 << XLB_BC_*; 0 = the kinds
 * XlbCollisionParams.bc_cells is meant for) into ``cells`` (linear indices (x*ny + y)*nz + z in no particular order, capacity ``capacity``
 * entries; sort them for coalesced access) and adds the number found to ``*count`` (device pointer, int64, zero it first; it may exceed ``capacity``, in which case only
 * the first ``capacity`` slots were written -- call again with a larger buffer).  cells == NULL only counts.  Setup-time
 * helper of the stepper (replaces the per-BC kernel specialisation of nse_stepper.py:351-380). */
int xlb_list_bc_cells(const XlbDomain* dom, const uint8_t* bc_mask, const XlbBcEntry* bc_table, int kind_mask, int x_begin, int x_end,
                      int64_t* cells, int64_t capacity, int64_t* count, void* stream);

/* Stream.__call__ (stream.py:29-62): f_out[l,x] = f_in[l, x - c_l] periodic. dtype = element type. */
int xlb_stream(const XlbDomain* dom, int dtype, const void* f_in, void* f_out, void* stream);
/* Macroscopic / ZeroMoment / FirstMoment (macroscopic.py:21-26, zero_moment.py:14-17, first_moment.py:14-18).
 * rho [1,N] and/or u [d,N] may be NULL (then not written); first moment divides by the computed rho. */
int xlb_macroscopic(const XlbDomain* dom, int f_dtype, const void* f, int out_dtype, void* rho, void* u, void* stream);
/* SecondMoment (second_moment.py:35-55): pi [d(d+1)/2, N] */
int xlb_second_moment(const XlbDomain* dom, int f_dtype, const void* f, int out_dtype, void* pi, void* stream);
/* QuadraticEquilibrium (quadratic_equilibrium.py:23-30): f[q,N] from rho[1,N], u[d,N] */
int xlb_equilibrium(const XlbDomain* dom, int in_dtype, const void* rho, const void* u, int f_dtype, void* f, void* stream);
/* BGK / KBC (bgk.py:27-32, kbc.py:40-79): fout = collide(f, feq, omega); all three share dtype */
int xlb_collide(const XlbDomain* dom, int collision, int dtype, const void* f, const void* feq, void* fout, double omega, void* stream);
/* SmagorinskyLESBGK as a stand-alone operator (smagorinsky_les_bgk.py:44-66); BGK / KBC as above; cp->has_force is not
 * applied here (ForcedCollision composes collide + macroscopic + equilibrium on the host side, forced_collision.py:51-57) */
int xlb_collide_ex(const XlbDomain* dom, int collision, const XlbCollisionParams* cp, int dtype, const void* f, const void* feq, void* fout,
                   double omega, void* stream);
/* BoundaryCondition.__call__ (boundary_condition.py:146-180 Warp kernel; bc_*.py JAX): applies BC ``id`` of
 * bc_table on cells with bc_mask == id, reading f_pre / f_post and writing f_out (may alias f_post).
 * assemble != 0 runs ExtrapolationOutflowBC.assemble_auxiliary_data (bc_extrapolation_outflow.py:106-133) instead. */
int xlb_apply_bc(const XlbDomain* dom, int dtype, const void* f_pre, const void* f_post, void* f_out, const uint8_t* bc_mask,
                 const uint8_t* missing_mask, const XlbBcEntry* bc_table, int id, int assemble, void* stream);

/* MomentumTransfer.__call__ (momentum_transfer.py:174-211, with FetchPopulations :63-70, stream-then-collide order):
 * force[a] += sum over cells x in planes [x_begin, x_end) with bc_mask == id and a non-solid centre, and over the missing
 * directions l, of c_opp(l)[a] * (f[opp l](x) + BC(stream(f))[l](x)).  ``force`` = 3 doubles on the device in KERNEL axes
 * (accumulated in FP64; zero it first), f = the post-collision populations the stepper returned. */
int xlb_momentum_transfer(const XlbDomain* dom, int dtype, const void* f_post_collision, const uint8_t* bc_mask, const uint8_t* missing_mask,
                          const XlbBcEntry* bc_table, int id, int x_begin, int x_end, double* force, void* stream);

/* IndicesBoundaryMasker (indices_boundary_masker.py:73-143), split in two device passes:
 * (1) tag: for every index column i (indices int32 [d][n], global coordinates minus ``origin``): cells outside
 *     the local block are dropped; bc_mask[cell] = id when id >= 0; solid[cell] = 1 when mark_solid != 0.
 *     Optional per-direction padding (pad != 0) tags index + c_l for all l (boundary_condition.py:123-136).
 * (2) missing: missing[l, x] = 1 iff x - c_l is outside the *global* domain or solid (the streamed mask, :134). */
int xlb_mask_tag(const XlbDomain* dom, const int32_t* indices, int64_t n, const int32_t origin[3], const int32_t global_shape[3], int id,
                 int mark_solid, int pad, uint8_t* bc_mask, uint8_t* solid, void* stream);
int xlb_mask_missing(const XlbDomain* dom, const uint8_t* solid, const int32_t origin[3], const int32_t global_shape[3],
                     uint8_t* missing_mask, void* stream);

/* missing_mask [q][N] (bool / uint8 planes, indices_boundary_masker.py:134-143) -> one uint32 per cell, bit l = missing[l] (SURVEY
 * 8(b) `xlb_pack_missing`; Appendix A: "a packed uint32 bitfield used by kernels").  Setup-time helper for
 * XlbCollisionParams.missing_bits; rerun it whenever missing_mask changes. */
int xlb_pack_missing(const XlbDomain* dom, const uint8_t* missing_mask, uint32_t* missing_bits, void* stream);

/* Allocation-free variant of the indices masker (xlb_mask_tag + xlb_mask_missing above).  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid allocation).  The two passes of the masker communicate through a per-cell ``solid`` scratch array
 * that the caller of xlb_mask_tag / xlb_mask_missing has to allocate; here that scratch is a file-scope
 * ``static __device__ __align__(256)`` array of XLB_MASK_STATIC_MAX_CELLS bytes inside the library (every GPU that loads the
 * library holds these 16 MiB until the process exits).  The caller still passes the inputs (index list) and the outputs
 * (bc_mask, missing_mask); the call itself allocates nothing -- no cudaMalloc / cudaFree, no allocating library routine.
 *   begin  != 0 : clear the scratch first (it keeps the previous call's contents otherwise) -- pass it with the first BC of a list;
 *   n > 0       : tag this BC exactly like xlb_mask_tag (same arguments), marking solid cells in the static scratch;
 *   finish != 0 : then run the xlb_mask_missing pass from the scratch into missing_mask -- pass it with the last BC of a list.
 * Blocks of more than XLB_MASK_STATIC_MAX_CELLS cells (nx*ny*nz) are rejected with XLB_EINVAL and a message.  All calls on a
 * device share the ONE array: calls that overlap on the GPU (different streams, or concurrent graph branches) race; serialise
 * them.  The array's address is resolved with cudaGetSymbolAddress once per device, with that device current. */
#define XLB_MASK_STATIC_MAX_CELLS (16 * 1024 * 1024)
int xlb_mask_indices_static(const XlbDomain* dom, const int32_t* indices, int64_t n, const int32_t origin[3], const int32_t global_shape[3],
                            int id, int mark_solid, int pad, int begin, int finish, uint8_t* bc_mask, uint8_t* missing_mask, void* stream);

/* MeshMaskerRay (boundary_masker/ray.py:43-77) + resolve_out_of_bound_kernel (mesh_boundary_masker.py:157-179): voxelises a triangle
 * soup (``triangles`` = float32 [n_triangles][3 vertices][3 coords], lattice units, GLOBAL kernel-axis coordinates, device
 * pointer).  Every cell of the local block from whose centre a lattice ray of length |c_l| hits the surface gets
 * bc_mask = id and missing_mask[opp l] = 1; cells tagged with ``id`` on the domain edge also get the directions that
 * pull from outside the global domain.  ``distance_keys`` (optional, uint64 [q][N], initialised to all ones) receives per
 * (direction, cell) the closest hit as (bits of t) << 32 | (bits of |hit - centre| / |c_l|), both FP32 (ray.py:72-76).
 * The BVH query of the reference (warp.mesh_query_ray) is replaced by one warp per triangle over the cells of its grown
 * bounding box with the watertight ray/triangle test (Woop, Benthin, Wald 2013); 3-D lattices only. */
int xlb_mask_mesh_ray(const XlbDomain* dom, const float* triangles, int64_t n_triangles, const int32_t origin[3],
                      const int32_t global_shape[3], int id, uint8_t* bc_mask, uint8_t* missing_mask, uint64_t* distance_keys, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* XLB_B200_H */
