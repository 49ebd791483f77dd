// pgt_kernels.cuh -- hand-written sm_100a kernels for the IK / thermo pair-sum path.
//
// Stage map (BASELINE.json north_star):
//   0/1  k_wrap_count (histogram + arrival rank), scan, k_scatter_ids, k_cell_sort_gather   (cell list, counting sort)
//   2/3  k_bonds_find<MODEL> + scan + k_bonds_fill_flat<MODEL>             (pair force + compacted half-bonds)
//   4/5  k_ik_quad (Cartesian grids, virial field), k_ik_gather2 (any point list, local fields), k_ik_adaptive, k_ik_elastic
//   6    k_thermo_pairs<MODEL>, k_thermo_final                             (global virial / energy / Born / kinetic)
// All arithmetic is FP64; no tensor cores (there is no dense contraction on this path).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "pgt_math.cuh"

#define PGT_MAX_GL 1024          // quadrature nodes kept in constant memory
#define PGT_NB_SLOTS 16          // force-carrying neighbours per particle kept by the single-search bond path
#define PGT_WRAP_MAX_ITERS 4096  // a position further than this many box lengths outside is a data error

// error bits raised by kernels (checked by the host after each stage)
enum { PGT_EFLAG_Z_NONZERO = 1, PGT_EFLAG_WRAP_RUNAWAY = 2, PGT_EFLAG_BOND_OVERFLOW = 4, PGT_EFLAG_QAG_FAILED = 8,
       PGT_EFLAG_QAG_CAPACITY = 16, PGT_EFLAG_BOUND = 32 /* a count exceeded the bound the host sized the frame's arrays for */ };

struct PgtBox {
    double lo[3], hi[3], L[3];
    int dim;
    int pstride;                 // components per row of the raw position array: 3 (x y z), or 2 (x y; pgt_frame_upload_xy)
};

struct PgtCells {
    int ncx, ncy;
    double inv_cwx, inv_cwy;     // 1 / cell width
    int reach;                   // cells to visit on each side for the max_rcut search
};

// Sorted (by cell, row-major) structure-of-arrays view of the live particles + half-bond CSR.
struct PgtFrameView {
    int n;                       // live particles (a bound when n_dev is set)
    const int* n_dev;            // device copy of the live count when the host did not wait for it (else NULL)
    const double* sx;            // wrapped x
    const double* sy;            // wrapped y
    const double2* sxy;          // (x, y) interleaved: 16-byte rows for the quad gather's bulk copies
    const double* sd;            // diameter
    const double* skxx;          // (m vx) vx   (NULL when the frame has no velocities)
    const double* skxy;          // (m vx) vy
    const double* skyy;          // (m vy) vy
    const int* sorig;            // original particle index
    const int* cell_start;       // ncx*ncy + 1
    const int* hb_off;           // n + 1  : half-bond CSR offsets
    // Half-bond i -> j records, structure of arrays (36 B per half-bond), everything the gather needs per
    // bond computed once instead of once per (grid point, bond):
    const double2* hb_b;         // b = r_j(image) - r_i, exactly as the reference's inner search forms it
    const double2* hb_cf;        // (|b|^2 * cscale,  f = -phi'(r)/r)
    const int* hb_j;             // sorted index of j
    const unsigned short* hb_pm; // per particle: bit k set when its k-th record is the bond's primary end (i < j); NULL when
                                 // the list was built by the two-search path (some particle has more than 16 bonds)
    long long n_hb;              // half-bonds in the list (host-side sizing of the walk gather's tiles)
    double hb_cscale;            // scale folded into hb_cf.x: 1/R^2 of the CG function the list was built for
};

struct PgtBondOut {              // destination of k_bonds<.,FILL>
    double2* b;
    double2* cf;
    int* j;
    long long cap;               // records the arrays hold
};

struct PgtGridView {
    int mode;                    // 0 cartesian, 1 point list
    int nx, ny;
    const double* xs;            // cartesian axes
    const double* ys;
    const double* pts;           // G x 3
    long long g0, g1;            // shard
    int monotone;                // cartesian: both axes non-decreasing (np.linspace)
    double span_x[4], span_y[4]; // cartesian: largest extent of an aligned tile of 8 w (x) / 4 w (y) grid points, w = index + 1; < 0 unknown
};

struct PgtIKArgs {
    PgtFrameView fr;
    PgtGridView grid;
    PgtBox box;
    PgtCells cells;
    PgtCG cg;
    double inv_rcutsq;           // 1 / R^2
    long long n_frame;           // particles of the frame as uploaded (before any kill or slab cull): sizes shard-independent capacities
    double max_rcut;             // the frozen neighbour-search radius
    int nq;                      // quadrature nodes (generic path)
    // 5-point rule in symmetric form on [0,1]: nodes 0.5 +- h_k, weights w_k (already times (b-a)/2)
    double gl5_h1, gl5_h1sq, gl5_h2, gl5_h2sq, gl5_w0, gl5_w1, gl5_w2;
    double gl_s[5], gl_w[5];     // the 5-point rule in GSL's visiting order (valid when nq == 5), for constant-bank operands
    const double* gl_nodes;      // per-context device table: nq nodes on [0,1] in GSL's visiting order
    const double* gl_weights;    //                           nq weights times (b-a)/2
    double* out_virial;          // (g1-g0) x 4
    double* out_kinetic;         // (g1-g0) x 4
    double* out_density;         // (g1-g0)
    unsigned obs_mask;
    double qag_epsabs, qag_epsrel;   // adaptive bond quadrature (GSL QAG): tolerances and interval limit
    int qag_limit;
};

struct PgtThermoOut { double v[32]; };   // virial 4, kinetic 4, energy 1, born 16 (2-D)

// ---- launch wrappers (defined in pgt_kernels.cu) ---------------------------------------------
struct PgtLaunchCounter { long long n; };

// In-place heapsort of a cell's particle ids (ascending): the path for cells with many members.
__device__ inline void pgt_heapsort_ids(int* a, int n)
{
    auto sift = [&](int root, int end) {
        const int v = a[root];
        for (;;) {
            int child = 2 * root + 1;
            if (child >= end) break;
            if (child + 1 < end && a[child + 1] > a[child]) ++child;
            if (a[child] <= v) break;
            a[root] = a[child];
            root = child;
        }
        a[root] = v;
    };
    for (int i = n / 2 - 1; i >= 0; --i) sift(i, n);
    for (int end = n - 1; end > 0; --end) {
        const int t = a[0]; a[0] = a[end]; a[end] = t;
        sift(0, end);
    }
}

// ---- Dim = 3 (pgt_dim3.cu): the reference's 3-D typedefs, parity kernels
struct Pgt3Cells {
    int ncx, ncy, ncz;
    double inv_cw[3];            // 1 / cell width per axis
};
struct Pgt3Frame {               // live particles sorted by cell ((cz ncy + cy) ncx + cx), ascending id inside a cell
    int n;
    const double *sx, *sy, *sz, *sd;
    const double *sm, *svx, *svy, *svz;   // NULL when the frame has no velocities
    const int* sorig;
    const int* cell_start;       // ncx ncy ncz + 1
};
struct Pgt3FrameOut { double *sx, *sy, *sz, *sd, *sm, *svx, *svy, *svz; };
struct Pgt3IKArgs {
    Pgt3Frame fr;
    PgtGridView grid;
    PgtBox box;
    Pgt3Cells cells;
    PgtCG cg;
    int nq;
    const double* gl_nodes;
    const double* gl_weights;
    int adaptive;
    double qag_epsabs, qag_epsrel;
    int qag_limit;
    double max_rcut;
    const double* disp;          // N x 3, original particle order (elastic field), or NULL
    double* out_virial;          // (g1 - g0) x 9, flat a + 3 b
    double* out_kinetic;         // (g1 - g0) x 9
    double* out_density;         // (g1 - g0)
    double* out_elastic;         // (g1 - g0) x 9
    unsigned obs_mask;
};
void pgt3_launch_wrap_count(const void* pos, int is_f32, long long N, PgtBox box, Pgt3Cells cells, int* cell_of, int* cell_slot,
                            int* cell_count, int* eflags, cudaStream_t st, PgtLaunchCounter* lc);
void pgt3_launch_sort_gather(const void* pos, const void* diam, const void* mass, const void* vel, int is_f32, PgtBox box, int* sorted_orig,
                             const int* cell_start, int ncells, Pgt3FrameOut out, cudaStream_t st, PgtLaunchCounter* lc);
cudaError_t pgt3_launch_ik(const PgtPot& pot, const Pgt3IKArgs& a, int* eflags, cudaStream_t st, PgtLaunchCounter* lc);
int pgt3_thermo_blocks(int n);
int pgt3_thermo_nv();
cudaError_t pgt3_launch_thermo(const PgtPot& pot, Pgt3Frame fr, PgtBox box, Pgt3Cells cells, double max_rcut, unsigned obs_mask,
                               const double* disp, double inv_vol, double* partials, int blocks, double* out_dev, int p_begin, int p_end,
                               cudaStream_t st, PgtLaunchCounter* lc);


// slab_on: drop particles whose wrapped x lies outside [slab_lo, slab_hi] under every periodic image (grid-point sharding)
void pgt_launch_wrap_count(const void* pos, int is_f32, long long N, PgtBox box, PgtCells cells, int* cell_of, int* cell_slot,
                           int* cell_count, int* eflags, int slab_on, double slab_lo, double slab_hi, cudaStream_t st,
                           PgtLaunchCounter* lc);
void pgt_launch_scan_exclusive(const int* in, int* out, long long n, int* block_sums, cudaStream_t st, PgtLaunchCounter* lc);
long long pgt_scan_scratch_ints(long long n);
void pgt_launch_scatter_ids(const int* cell_of, const int* cell_slot, long long N, const int* cell_start, int* sorted_orig,
                            cudaStream_t st, PgtLaunchCounter* lc);
// per cell: members ordered by original index, then gathered (promoted, wrapped) into the sorted structure of arrays
void pgt_launch_cell_sort_gather(const void* pos, const void* diam, const void* mass, const void* vel, int is_f32, PgtBox box,
                                 int* sorted_orig, const int* cell_start, int ncells, int n_cap, double* sx, double* sy, double2* sxy,
                                 double* sd, double* skxx, double* skxy, double* skyy, cudaStream_t st, PgtLaunchCounter* lc);
// mode 0: count force-carrying half-bonds per particle into counts[n+1]; mode 1: fill the CSR arrays;
// mode 2/3: same for plain neighbours within max_rcut (export only);
// mode 4: count AND keep the neighbours (nbr_j = [PGT_NB_SLOTS][n] scratch, out.j = error-flag word, out.b = the
//         per-particle primary masks, unsigned short[n]);
// mode 5: fill the CSR arrays from that scratch (counts = scanned offsets), warp-flattened; mode 6: same, thread per particle.
void pgt_launch_bonds(int mode, const PgtPot& pot, PgtFrameView fr, PgtBox box, PgtCells cells, double max_rcut, int* counts,
                      PgtBondOut out, double cscale, int* nbr_j, cudaStream_t st, PgtLaunchCounter* lc);
// rewrites hb_cf.x of every record for a new scale (the CG radius changed after the list was built)
void pgt_launch_rescale_bonds(const double2* hb_b, double2* hb_cf, long long n_hb, double cscale, cudaStream_t st,
                              PgtLaunchCounter* lc);
// variant 0: automatic (quad gather on Cartesian grids with the virial field only, else two-phase); 2: two-phase kernel;
// 7: quad gather where applicable
cudaError_t pgt_launch_ik(const PgtIKArgs& a, int variant, cudaStream_t st, PgtLaunchCounter* lc);
// virial field with the adaptive (QAG, 15-point Gauss-Kronrod) bond quadrature; failures are OR-ed into *eflags
cudaError_t pgt_launch_ik_adaptive(const PgtIKArgs& a, int* eflags, cudaStream_t st, PgtLaunchCounter* lc);
// [p_begin, p_end): the sorted particles whose pairs (orig_i < orig_j, visited from i) this launch sums
void pgt_launch_thermo(const PgtPot& pot, PgtFrameView fr, PgtBox box, PgtCells cells, double max_rcut, unsigned obs_mask,
                       double inv_vol, double* partials, int max_blocks, double* out_dev, int p_begin, int p_end, cudaStream_t st,
                       PgtLaunchCounter* lc);
int pgt_thermo_blocks(int n);
// ElasticVirialStress (needs per-particle displacements, rows of the original particle order): global sum and field
void pgt_launch_thermo_elastic(const PgtPot& pot, PgtFrameView fr, PgtBox box, PgtCells cells, double max_rcut, const double* disp,
                               double inv_vol, double* partials, int blocks, double* out_dev, cudaStream_t st, PgtLaunchCounter* lc);
cudaError_t pgt_launch_ik_elastic(const PgtPot& pot, const PgtIKArgs& a, const double* disp, double max_rcut, int adaptive, double* out,
                                  int* eflags, cudaStream_t st, PgtLaunchCounter* lc);
void pgt_launch_ik_count(const PgtIKArgs& a, unsigned long long* out3, cudaStream_t st, PgtLaunchCounter* lc);
// count + flags of a stage to mapped host memory, flags and bound violations OR-ed into a sticky device word
void pgt_launch_publish_count(const int* count, const int* flags, long long bound, int* sticky, int* dst_mapped, int* dst_sticky,
                              cudaStream_t st, PgtLaunchCounter* lc);
// copies nwords0 + nwords1 (<= 64) 32-bit words to mapped host memory
void pgt_launch_publish(const void* s0, int nwords0, const void* s1, int nwords1, void* dst_mapped, cudaStream_t st, PgtLaunchCounter* lc);
// (xx, xy, yx, yy) per point -> (xx, xy, yy) per point
void pgt_launch_pack_sym(const double* in4, double* out3, long long n, cudaStream_t st, PgtLaunchCounter* lc);
void pgt_launch_fp64_probe(double* out, int iters, int blocks, int threads, cudaStream_t st, PgtLaunchCounter* lc);
