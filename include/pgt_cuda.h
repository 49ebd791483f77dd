/* pgt_cuda.h -- C-ABI of libpgt_cuda.so: the B200 (sm_100a) implementation of PyGlassTools'
 * Irving-Kirkwood / thermo pair-sum hot path.
 *
 * Plain C, opaque handle, plain pointers and sizes; no C++/torch types cross this boundary.
 * Every entry point returns 0 on success or a negative pgt_status; pgt_last_error(ctx) gives
 * the message.  The caller owns every host buffer, the library owns every device buffer.
 * One pgt_ctx drives ONE device on ONE stream; multi-GPU = one ctx (and normally one process)
 * per device, each given a shard [g0,g1) of the grid points (pgt_grid_set_shard).
 *
 * Each entry point names the reference interface (file:line under /root/reference/pyglasstools/)
 * whose role it takes over.  There is no CPU fallback behind this header: without a CUDA device
 * pgt_ctx_create fails.
 */
#ifndef PGT_CUDA_H
#define PGT_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pgt_ctx pgt_ctx;

typedef enum {
    PGT_OK = 0,
    PGT_ERR_CUDA = -1,        /* a CUDA runtime call or kernel failed */
    PGT_ERR_ARG = -2,         /* invalid argument / call order */
    PGT_ERR_UNSUPPORTED = -3, /* valid in the reference, not implemented on device (3-D, ...) */
    PGT_ERR_STATE = -4,       /* a required earlier stage has not run */
    PGT_ERR_DATA = -5,        /* the frame violates a documented precondition (box too small, z != 0 in 2-D, ...) */
    PGT_ERR_NCCL = -6
} pgt_status;

/* pair-model ids: the eight classes exported by potential/module-potential.cpp:13-34 */
enum { PGT_POT_LJ = 0, PGT_POT_FORCESHIFTED_LJ = 1, PGT_POT_POLY12 = 2, PGT_POT_POLY18 = 3, PGT_POT_POLY10 = 4,
       PGT_POT_POLY106 = 5, PGT_POT_POLYYUKAWA = 6, PGT_POT_POLYLJ = 7 };
/* coarse-graining kernels: irvingkirkwood/cgfunc/module-cgfunc.cpp:12-19 */
enum { PGT_CG_OCTIC = 0, PGT_CG_RECT = 1, PGT_CG_MOLLIFIER = 2 };
/* quadrature modes: FixedPointCGFunc (CoarseGrainFunction.h:56-92); AdaptiveCGFunc (:94-134) via pgt_cg_set_adaptive */
enum { PGT_QUAD_FIXEDPOINT = 0, PGT_QUAD_ADAPTIVE = 1 };
/* observable masks */
enum { PGT_OBS_VIRIAL = 1u, PGT_OBS_KINETIC = 2u, PGT_OBS_DENSITY = 4u,      /* IK fields */
       PGT_OBS_ENERGY = 8u, PGT_OBS_BORN = 16u,                             /* thermo only */
       PGT_OBS_ELASTIC = 32u };                                             /* IK field (needs displacements) */
/* observable ids for downloads */
enum { PGT_FIELD_VIRIAL = 0, PGT_FIELD_KINETIC = 1, PGT_FIELD_DENSITY = 2, PGT_FIELD_ELASTIC = 3 };
/* stage ids for pgt_stage_time_ms */
enum { PGT_STAGE_UPLOAD = 0, PGT_STAGE_CELLS = 1, PGT_STAGE_BONDS = 2, PGT_STAGE_IK = 3, PGT_STAGE_THERMO = 4,
       PGT_STAGE_DOWNLOAD = 5, PGT_STAGE_COUNT = 6 };

/* ---- context ------------------------------------------------------------------------------- */
/* Replaces: MPI::Communicator construction at import (module-pyglasstools.cpp:50-58).          */
int pgt_ctx_create(int device, pgt_ctx** out);
int pgt_ctx_destroy(pgt_ctx* ctx);
const char* pgt_last_error(const pgt_ctx* ctx);      /* ctx may be NULL: last create error */
int pgt_device_count(int* n);
int pgt_synchronize(pgt_ctx* ctx);
/* The CUDA stream (cudaStream_t) every kernel of this ctx is launched on, for event timing. */
void* pgt_stream(pgt_ctx* ctx);

/* ---- models -------------------------------------------------------------------------------- */
/* Replaces: ShortRangePairPotential<M>(scaled_rcut, params) (potential/PairPotential.h:69-123).
 * params: the vector the Python factories build (potential/__init__.py:78-143).               */
int pgt_potential_set(pgt_ctx* ctx, int model_id, double scaled_rcut, const double* params, int nparams);
/* Replaces: FixedPointCGFunc<D>(order, rcut) (cgfunc/CoarseGrainFunction.h:56-92) and the GSL
 * glfixed table (cgfunc/FixedPointQuadrature.h:35).                                            */
int pgt_cg_set(pgt_ctx* ctx, int kernel_id, double cg_rcut, int quad_mode, int order);
/* Replaces: AdaptiveCGFunc<D>(maxiter, relerr, abserr, rcut) (cgfunc/CoarseGrainFunction.h:94-134) and its
 * gsl_integration_qag(.., epsabs, epsrel, limit, key = 1, ..) call (cgfunc/Quadrature.h:39-51): global adaptive
 * bisection with the 15-point Gauss-Kronrod pair.  A bond on which GSL would exit through its error handler (the
 * reference aborts there) makes pgt_ik_compute return PGT_ERR_DATA.                                              */
int pgt_cg_set_adaptive(pgt_ctx* ctx, int kernel_id, double cg_rcut, double epsabs, double epsrel, int limit);
/* Replaces: Calculator::max_rcut, frozen at construction (Calculator.h:37-46).                 */
int pgt_set_max_rcut(pgt_ctx* ctx, double max_rcut);
/* Optional tuning: target cell side of the device cell list (default max_rcut/2).              */
int pgt_set_cell_size(pgt_ctx* ctx, double cell_side);

/* ---- frame --------------------------------------------------------------------------------- */
/* Replaces: ParticleSystem ctor + setParticleSystemData + init_neighbour_search
 * (ParticleSystem.cc:3-30; Aboria NeighbourSearchBase.h:208-237) and SimBox (SimBox.cc:25-52).
 * pos: N x 3 (float or double), diam: N, mass: N or NULL (=1), vel: N x 3 or NULL (=0).
 * box: Lx,Ly,Lz (Lz forced to 1 when dim==2 as utils.py:15-16 does).  Host pointers; the copy
 * is asynchronous on the ctx stream from the caller's (ideally pinned) memory.
 * dim == 3 selects the reference's Dim = 3 typedefs (module-irvingkirkwood.cpp:9-13, module-thermo.cpp:8-20): the box is
 * periodic on all three axes (SimBox.cc:33), fields are Dim*Dim = 9 doubles per grid point, pgt_thermo_compute returns
 * 9 + 9 + 1 + 81 doubles, pgt_thermo_elastic 9.  The 3-D stages are parity kernels (the reference's nested searches, no
 * bond records: pgt_build_bonds is a no-op); the parity exports, pgt_ik_count_work, pgt_ik_download_sym_async and
 * pgt_thermo_compute_static serve 2-D frames only (PGT_ERR_UNSUPPORTED otherwise).                                  */
int pgt_frame_upload(pgt_ctx* ctx, const void* pos, const void* diam, const void* mass, const void* vel,
                     int is_f32, int64_t N, const double box[3], int dim);
/* The same for a 2-D frame whose positions the caller holds as N x 2 rows (x, y): z = 0 for every particle, 12 instead of
 * 16 bytes per float32 particle cross the host link (25 % of the upload).  What ParticleSystem's ctor would receive with
 * a zero third column (ParticleSystem.cc:3-30); same stages, same bits as pgt_frame_upload of the N x 3 array.           */
int pgt_frame_upload_xy(pgt_ctx* ctx, const void* pos_xy, const void* diam, const void* mass, const void* vel,
                        int is_f32, int64_t N, const double box[3]);
/* One rank's PART of a frame (grid-point sharding over GPUs: every GPU needs the whole frame, but each rank pushes
 * only 1/nranks of it over PCIe and pgt_comm_allgather_frame assembles it on every GPU over NVLink).  Parts are
 * ceil(N / nranks) particles each, rank r's starting at particle r * ceil(N / nranks) (the last may be short):
 * pos_part / diam_part point at the first particle of THIS rank's part, `first` and `count` name it.  float32 or
 * float64 positions and diameters only (no mass / velocity).  Replaces the per-rank GSD read of the whole frame
 * (utils.py:97-121 under MPI: every rank reads every frame).                                                       */
int pgt_frame_upload_part(pgt_ctx* ctx, const void* pos_part, const void* diam_part, int is_f32, int64_t first, int64_t count,
                          int64_t N, const double box[3], int dim);
/* Same, from DEVICE pointers already resident in HBM on ctx's device (no host copy).           */
int pgt_frame_set_device(pgt_ctx* ctx, const void* d_pos, const void* d_diam, const void* d_mass, const void* d_vel,
                         int is_f32, int64_t N, const double box[3], int dim);
/* Stage 1: wrap/kill + counting-sort cell list.  Replaces Aboria CellList build (CellList.h:143-194). */
int pgt_build_cells(pgt_ctx* ctx);
/* Stages 2-3: pair forces + per-particle compacted half-bond list.  Replaces the inner
 * euclidean_search + getPairForceDivR of IrvingKirkwood.h:93-105.                              */
int pgt_build_bonds(pgt_ctx* ctx);
int pgt_frame_counts(pgt_ctx* ctx, int64_t* n_live, int64_t* n_halfbonds, int64_t* n_cells);

/* ---- grid ---------------------------------------------------------------------------------- */
/* Replaces: ikcalculator grid construction + scatter_v (irvingkirkwood/__init__.py:45-63).
 * Cartesian: point g = i*ny + j at (xs[i], ys[j], 0), i-major.                                 */
int pgt_grid_set_cartesian(pgt_ctx* ctx, const double* xs, int nx, const double* ys, int ny);
/* Arbitrary points (radialcalculator, irvingkirkwood/__init__.py:80-166): xyz is G x 3.        */
int pgt_grid_set_points(pgt_ctx* ctx, const double* xyz, int64_t G);
/* This ctx computes global points [g0,g1) (np.array_split boundaries are the caller's job).  On a Cartesian grid with
 * non-decreasing axes a strict sub-range also restricts stages 1-3 to the particles the shard's tiles can reach
 * (option "slab_cull", default 1): the field is bit-identical, pgt_frame_counts reports the particles kept, and
 * pgt_thermo_compute refuses such a frame (build it with the whole grid, or set "slab_cull" to 0).                 */
int pgt_grid_set_shard(pgt_ctx* ctx, int64_t g0, int64_t g1);

/* ---- compute ------------------------------------------------------------------------------- */
/* Stages 4-5.  Replaces IrvingKirkwood::compute (irvingkirkwood/IrvingKirkwood.h:74-110).      */
int pgt_ik_compute(pgt_ctx* ctx, unsigned obs_mask);
/* Copies a field of the local shard to host: (g1-g0) x Dim*Dim doubles, col-major per point
 * (index a + Dim*b, CoarseGrainedField.h:160-183), or (g1-g0) doubles for the density.         */
int pgt_ik_download(pgt_ctx* ctx, int field_id, double* out);
/* Same copy, enqueued on the ctx stream without waiting: `out` (page-locked, see pgt_host_alloc) is valid after
 * pgt_synchronize(ctx).  With two contexts used alternately, frame k+1's upload and compute overlap frame k's
 * read-back (PCIe is full duplex) -- the trajectory loop of pyglasstools.analyze, double-buffered.              */
int pgt_ik_download_async(pgt_ctx* ctx, int field_id, double* out);
/* Same for a 2x2 tensor field, 3 doubles per grid point (xx, xy, yy): the stress tensors are symmetric by construction
 * (VirialStress::compute, ListofCommonObservables.h:49-70: F (x) r with F parallel to r), the copy is 25 % smaller. */
int pgt_ik_download_sym_async(pgt_ctx* ctx, int field_id, double* out);
/* Waits for the read-backs enqueued by pgt_ik_download_async only (not for kernels of later frames).        */
int pgt_ik_download_wait(pgt_ctx* ctx);
/* Waits until at most `keep` (0..3) of the asynchronous read-backs enqueued so far are still in flight.  With keep + 1
 * host buffers used in turn the caller consumes frame s - keep while later frames are on the device and frame s + 1 is
 * being uploaded -- one context, one host thread, the trajectory loop of pyglasstools.analyze fully pipelined (uploads,
 * stages and read-backs run on three streams of the context; frames whose counts are deferred never make the host wait
 * in between).                                                                                                       */
int pgt_ik_download_wait_keep(pgt_ctx* ctx, int keep);
/* Device pointer of a field of the local shard (for NCCL / zero-copy consumers).               */
int pgt_ik_device_ptr(pgt_ctx* ctx, int field_id, void** d_ptr, int64_t* n_doubles);
/* Stage 6.  Replaces ThermoCalculator::compute + divideByVolume (thermo/ThermoCalculator.h:58-90).
 * out layout (doubles): [0..Dim^2) virial, [Dim^2..2Dim^2) kinetic, [2Dim^2] potential energy,
 * then Dim^4 Born tensor (col-major flat); all already multiplied by 1/vol.  Entries whose
 * observable is not in obs_mask are 0.  out must hold 2*Dim^2 + 1 + Dim^4 doubles.              */
int pgt_thermo_compute(pgt_ctx* ctx, unsigned obs_mask, double* out);
/* The same without a host synchronisation: `out` (25 doubles, page-locked for a truly asynchronous copy) is valid after
 * pgt_synchronize.  With the option "async_counts" a trajectory loop of pgt_frame_upload / pgt_build_cells /
 * pgt_thermo_compute_async never waits for the device (the reference's loop, thermo/ThermoCalculator.h:68-90, is
 * synchronous: this is the thermo counterpart of pgt_ik_download_async).  2-D frames.                              */
int pgt_thermo_compute_async(pgt_ctx* ctx, unsigned obs_mask, double* out);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same inputs, same kernels, same
 * block partition and same `out` layout as pgt_thermo_compute, but the per-block partial sums and the reduced result
 * live in one file-scope `static __device__` array of the library (PGT_THERMO_STATIC_MAX_BLOCKS * 25 + 32 doubles =
 * 473 856 bytes, held by every GPU that loads the library until the process exits) instead of the context's device
 * buffers: the call makes no cudaMalloc, cudaFree or other allocating call.  Frames with more than
 * PGT_THERMO_STATIC_MAX_PARTICLES live particles are refused with PGT_ERR_ARG (use pgt_thermo_compute).  All contexts
 * and all calls on a device share the array, which keeps the last call's contents (the call clears what it needs);
 * calls that overlap on the GPU (two contexts = two streams on one device) race: serialise them.                    */
#define PGT_THERMO_STATIC_MAX_BLOCKS 2368
#define PGT_THERMO_STATIC_MAX_PARTICLES 2097152
int pgt_thermo_compute_static(pgt_ctx* ctx, unsigned obs_mask, double* out);
/* Particle-sharded thermo sums over GPUs: this ctx sums the pairs (id_i < id_j) visited from part `part` of `nparts`
 * equal shares of the sorted particles; the nparts partial results add up to pgt_thermo_compute's (sum them with
 * pgt_comm_allreduce_sum: the role of MPI reductions the reference never needed, ThermoCalculator being serial).
 * (0, 1) restores the whole frame.                                                                                  */
int pgt_thermo_set_shard(pgt_ctx* ctx, int part, int nparts);
/* Replaces: ParticleSystem::setDisplacement (ParticleSystem.cc:60-67): N x 3 doubles, one row per particle of the frame
 * last uploaded, in the frame's own order.  A new frame starts with zero displacements (ParticleSystem.cc:28).   */
int pgt_frame_set_displacement(pgt_ctx* ctx, const double* disp_xyz, int64_t N);
/* Replaces: GlobalElasticVirialStress (ListofCommonObservables.h:86-123 through ThermoCalculator::compute): the
 * Dim^2 sums of F1 (x) (u+r) - F (x) r over pairs id_i < id_j within max_rcut, times 1/vol.  The field version is
 * pgt_ik_compute with PGT_OBS_ELASTIC, read back as PGT_FIELD_ELASTIC.                                            */
int pgt_thermo_elastic(pgt_ctx* ctx, double* out /* Dim^2 */);

/* ---- parity exports / instrumentation ------------------------------------------------------ */
/* Ordered force-carrying pairs (orig_i, orig_j) and r_ij = r_i - r_j(image) as the reference's
 * inner loop sees them; call with cap=0 to get the count.                                      */
int pgt_export_bonds(pgt_ctx* ctx, int64_t cap, int64_t* n, int32_t* i, int32_t* j, double* rij_xy, double* fdivr);
/* Ordered neighbour pairs within max_rcut (inclusive), force-carrying or not.                  */
int pgt_export_neighbors(pgt_ctx* ctx, int64_t cap, int64_t* n, int32_t* i, int32_t* j);
/* Wrapped positions in original order (x,y per particle) and alive flags.                      */
int pgt_export_wrapped(pgt_ctx* ctx, double* xy, uint8_t* alive);
/* Device time of the last execution of a stage (CUDA events on the ctx stream).                */
int pgt_stage_time_ms(pgt_ctx* ctx, int stage, float* ms);
/* Number of kernels launched by this ctx since creation.                                       */
int64_t pgt_kernel_launches(const pgt_ctx* ctx);
/* Scalar host-side evaluation of the pair model (getPairEnergy/ForceDivR/BondStiffness/getRcut;
 * PairPotential.h:92-122) and CG kernel; what: 0 energy, 1 force/r, 2 stiffness, 3 rcut^2.     */
double pgt_pair_eval(int model_id, double scaled_rcut, const double* params, int nparams,
                     const double rij[3], double di, double dj, int what);
double pgt_cg_delta(int kernel_id, double cg_rcut, double dr_sq);
double pgt_cg_bond(int kernel_id, double cg_rcut, int order, const double x[3], const double ri[3], const double rij[3]);
/* AdaptiveCGFunc::getBondFunc (CoarseGrainFunction.h:123-126); *status (may be NULL) receives 0 or the QAG exit code */
double pgt_cg_bond_adaptive(int kernel_id, double cg_rcut, double epsabs, double epsrel, int limit, const double x[3],
                            const double ri[3], const double rij[3], int* status);
/* FP64 FMA peak probe: runs a dependent-chain DFMA kernel, returns achieved TFLOP/s.            */
int pgt_fp64_peak_probe(pgt_ctx* ctx, double* tflops);
/* Algorithmic work of the IK gather: sum over the local shard of U(x), the
 * number of unique force-carrying bonds with at least one endpoint within cg_rcut of x; also the
 * number of (grid point, half-bond) items the gather kernel actually integrates.                */
int pgt_ik_count_work(pgt_ctx* ctx, int64_t* sum_unique_bonds, int64_t* sum_halfbond_items, int64_t* sum_particles);
/* User events on the ctx stream (slots 0..15) for timing several stages as one region.          */
int pgt_event_record(pgt_ctx* ctx, int slot);
int pgt_event_elapsed_ms(pgt_ctx* ctx, int slot_begin, int slot_end, float* ms);
/* Page-locked host memory for the caller's frame / field buffers (cudaHostAlloc / cudaFreeHost). */
int pgt_host_alloc(void** ptr, uint64_t bytes);
int pgt_host_free(void* ptr);
/* Page-locks / releases an EXISTING host allocation (cudaHostRegister): copies from and to it then run at link speed and
 * asynchronously.  No counterpart in the reference.  PGT_ERR_CUDA when the driver refuses: the buffer stays usable, pageable.
 * Unregister before the memory is freed.                                                                              */
int pgt_host_register(void* ptr, uint64_t bytes);
int pgt_host_unregister(void* ptr);
/* Tuning knobs: "ik_variant" (0 automatic, 2 two-phase gather, 7 quad gather where it applies);
 * "async_counts" (1 default: a frame of the same shape as the last one counted sizes its arrays from that frame's particle
 *   and half-bond counts plus a margin and does not wait for its own; counts and errors reach the host at the next
 *   pgt_synchronize / pgt_ik_download / pgt_ik_download_wait / pgt_frame_counts; 0: every stage waits for its count);
 * "slab_cull" (1 default: a grid shard restricts stages 1-3 to the particles it can reach, 0: every rank builds the whole frame);
 * "bond_single_search" (1 default: one neighbour search per frame, 0: separate count and fill searches);
 * "bond_fill_flat" (1 default: warp-flattened record writer, 0: thread per particle).  Every setting gives the same
 * bits.  Environment, read once: PGT_IK_VARIANT, PGT_BOND_FILL_FLAT, PGT_SCAN_SINGLE_PASS (0: three-kernel scan). */
int pgt_set_option(pgt_ctx* ctx, const char* name, double value);

/* ---- multi-GPU (NCCL over NVLink) ---------------------------------------------------------- */
/* 128-byte NCCL unique id, to be broadcast by the caller's own plumbing.                       */
int pgt_comm_unique_id(void* id128);
int pgt_comm_init(pgt_ctx* ctx, int nranks, int rank, const void* id128);
/* Final field gather: every rank contributes its shard, out_dev (device, G_total*ncomp doubles
 * on every rank) receives all shards in rank order.  counts[r] = doubles contributed by rank r. */
int pgt_comm_allgather_field(pgt_ctx* ctx, int field_id, const int64_t* counts, void* out_dev);
/* Assembles the frame on every GPU from the parts pushed by pgt_frame_upload_part: one ncclAllGather per array on the
 * ctx stream (NVLink / NVSwitch), in place in the frame's device buffers.  No host synchronisation.              */
int pgt_comm_allgather_frame(pgt_ctx* ctx);
/* Sum-reduce n doubles in place on the host buffer via the device (thermo partial sums).       */
int pgt_comm_allreduce_sum(pgt_ctx* ctx, double* inout_host, int n);
int pgt_comm_destroy(pgt_ctx* ctx);

#ifdef __cplusplus
}
#endif
#endif /* PGT_CUDA_H */
