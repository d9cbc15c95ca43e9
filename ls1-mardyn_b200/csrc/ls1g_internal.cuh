// ls1g_internal.cuh — shared declarations of the B200 (sm_100a) short-range force + integrate path.
//
// One ls1g_ctx = one ParticleContainer (reference: src/particleContainer/LinkedCells.h) resident in
// HBM as a cell-sorted SoA.  Kernels live in the .cu files next to this header; the C ABI
// (include/ls1g.h) is implemented in api.cu.  No CPU fallback exists anywhere in this directory.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>
#include <string.h>
#include <string>
#include <vector>

#include "../../include/ls1g.h"

// ------------------------------------------------------------------------------------------------
// Grid geometry (LinkedCells::rebuild / constructor, src/particleContainer/LinkedCells.cpp:42-204 and
// CellBorderAndFlagManager.h:99-157).  Passed BY VALUE to kernels (lives in the kernel-parameter
// constant bank).
// ------------------------------------------------------------------------------------------------
static inline long long host_double_bits(double x) { long long b; memcpy(&b, &x, sizeof b); return b; }

struct Grid {
	double box_min[3], box_max[3];
	double halo_min[3], halo_max[3];
	double cell_len[3], cell_len_recip[3];
	double box_len[3];        // box_max - box_min : periodic shift magnitude (DomainDecompBase.cpp:175)
	double cutoff;            // halo slab width (DomainDecompBase.cpp:293-348)
	int    cpd[3];            // cells per dimension incl. one halo layer on each side
	int    n_cells;
	int    self_periodic[3];  // 1: this rank is its own neighbour in that dimension (rank_grid[d]==1)
};

// Per-component constants (src/molecules/Component.h, FullMolecule::updateMassInertia FullMolecule.h:88-101)
struct CompDev {
	double mass, inv_mass_unused;
	double I[3], invI[3];
	int    rot_dof;
	int    n_lj, n_ch, n_di, n_qu;
	int    lj_id0;            // Component::getLookUpId (EnsembleBase.cpp:91-103): global id of LJ site 0
	int    site0;             // offset of this component's first site row in the body-frame site table
	int    n_sites;           // n_lj + n_ch + n_di + n_qu
};

// Body-frame site row (Site.h:84-300).  type: 0 LJ, 1 charge, 2 dipole, 3 quadrupole.
struct SiteDev {
	double d[3];              // body-frame offset
	double e[3];              // body-frame orientation (dipole / quadrupole)
	double val;               // q | |mu| | |Q| ; unused for LJ
	int    type;
	int    ljid;              // global LJ id (LJ sites)
};

enum { SITE_LJ = 0, SITE_CH = 1, SITE_DI = 2, SITE_QU = 3 };

// ------------------------------------------------------------------------------------------------
// Molecule storage.  All arrays have `cap` entries, index = slot in cell-sorted order.
// pos.w carries the component id (low 32 bits of the bit pattern) so that one 32-byte load gives a
// neighbour's centre AND its component.
// ------------------------------------------------------------------------------------------------
struct MolBuf {
	double4*  pos;    // x y z | cid bits
	double*   vel;    // [cap][3]
	double4*  q;      // w x y z           (only if ctx->has_rot)
	double*   L;      // [cap][3]          (only if ctx->has_rot)
	uint64_t* id;
	uint32_t* cell;   // linear cell index | HALO_FLAG
};

static const uint32_t HALO_FLAG = 0x80000000u;

struct Timer { double ms; uint64_t launches; };
struct PendingEvt { cudaEvent_t a, b; int which; int launches; };

struct ls1g_ctx {
	ls1g_config cfg;
	Grid grid;                              // the reference's cell grid (LinkedCells::rebuild): cell queries, C08 dump, download
	Grid bgrid;                             // the grid the device sort bins on: == grid, except inside the neighbour-list loop of
	                                        // ls1g_run where its cells are >= cutoff + skin (lists.cu)
	cudaStream_t stream;

	// components
	int n_comp, n_site_rows, S_lj;          // S_lj = total LJ sites over all components
	int max_sites;                          // max sites of any component (site-cache stride)
	int has_rot;                            // any component needs q/L (multi-site or rot_dof>0)
	int simple_1clj;                        // exactly one component: one LJ site at the origin, nothing else
	int force_variant;                      // 0 = default kernels; 1 = v1 thread-per-molecule kernels (env LS1G_FORCE_KERNEL=simple)
	std::vector<CompDev> comps_h;
	std::vector<SiteDev> sites_h;
	std::vector<double>  eps24_h, sig2_h, shift6_h;
	CompDev* comps; SiteDev* sites;         // device copies
	double *eps24, *sig2, *shift6;          // device [S_lj*S_lj]
	double epsRFInvrc3, rc2, rclj2;

	// molecules
	size_t cap;                             // allocated slots
	size_t n_slots;                         // slots in use (owned + halo copies), cell-sorted if sorted
	size_t n_owned, n_halo;
	size_t n_uploaded;                      // n of the last ls1g_upload (ls1g_upload_forces refers to it)
	uint32_t max_row;                       // owned molecules of the fullest cell row after the last sort
	uint32_t max_bundle;                    // ... of the fullest 2 x 2 bundle of cell rows
	int    sorted;                          // cell[] / cell_start valid
	int    forces_valid;
	MolBuf cur, alt;                        // double buffer for the sort
	double *F, *M, *Vi;                     // [cap][3]; M only if has_rot
	double *F_alt, *M_alt, *Vi_alt;                // targets when forces are carried through a re-sort
	double4* site_cache;                    // [cap*cache_stride] world-frame site data (multi-site path)
	int     cache_stride;
	void*   stage; size_t stage_bytes;
	// neighbour-list mode of ls1g_run (lists.cu): frozen slot layout between rebuilds
	double   skin;                          // list radius = cutoff + skin; 0 disables the list mode
	int      in_list_mode;                  // the slots are in the frozen neighbour-list layout (bin grid, lists valid or expired)
	int      list_disabled;                 // set when a buffer limit was hit once: stay on the rebuild-every-step path
	uint32_t *halo_src; uint8_t* halo_code; // [cap] per halo slot: slot of its source molecule, periodic image code
	uint32_t *slot_dst;                     // [cap] final slot of scatter entry s (sort scratch)
	double4* pos_build;                     // [cap] positions when the lists were built (displacement criterion)
	unsigned short* lists; size_t lists_cap;// [blocks][LCAP][128] tile indices
	uint32_t* lcount; size_t lcount_cap;    // [blocks*128]
	void* plans;                            // [blocks] NLPlan of every block of the last build (lists.cuh), sized with bmap
	uint32_t* bmap; unsigned char* bflag; size_t bmap_cap;   // boundary blocks of the last build (lists.cuh: NLArgs::bmap)
	uint32_t n_boundary_blocks;
	int      force_span;                    // the two launches of an overlapped step run side by side: time them as one interval (lists.cu)
	cudaStream_t stream_comm;               // high-priority stream of the overlapped halo refresh (multi-rank list steps)
	cudaEvent_t ev_pos, ev_halo;
	int      track_disp;                    // integrate kernels track max |r - pos_build|^2 into d_flags[0]
	uint64_t list_rebuilds;                 // list builds since the context was created (ls1g_list_stats)
	int      list_R;                        // cell rows per block of the list kernels: 4 (2 x 2 bundle) or 1 (lists.cuh)
	int      list_multisite;                // rigid multi-site systems take the list loop too (LS1G_LIST_MULTISITE=0: per-step cell scan)
	int      list_permute;                  // bank-aware list order: 1 in the build kernel (default), 0 none, 2 separate whole-list pass (LS1G_LIST_PERMUTE)
	uint64_t list_steps;                    // list steps completed since the context was created
	uint64_t list_steps_at_build, list_life;   // ... at the last build; lifetime (steps) of the lists that build replaced
	int      list_zchk;                     // the last build saw a (nearly) coincident pair: force kernels keep the r2 == 0 test
	int      list_tcap;                     // tile capacity used by the list force kernel (max tile of the last build, padded)
	int      list_bpr;                      // blocks per cell row
	int      build_tcap;                    // 9-row tile capacity of the build kernel, raised once on overflow
	int      list_lcap;                     // list entries per molecule (multiple of 4), grown on overflow
	unsigned long long* d_flags;            // [16]: 0 max displacement^2 since the BUILD (bits) | 1 the same since the last PRUNE | 2 max tile |
	                                        //       3 guard word of the next step: bits of 0.0 = go, 2.0 = lists expired (nl_guard_kernel) |
	                                        //       5 boundary blocks | 6 near-pair flag of the last build | 7 overflow bits of the last build (NL_F_OVF)
	// inner lists (lists.cuh: nl_prune_kernel): the outer lists (radius cutoff + skin) are pruned every few steps to cutoff + skin_in;
	// the force kernels read the pruned copy.  1CLJ only; skin_in = 0 disables them.
	double   skin_in;                       // inner skin
	int      skin_user;                     // LS1G_SKIN given: the outer skin is not adapted to the kind of system
	unsigned short* lists_in; size_t lists_in_cap;
	uint32_t* lcount_in; size_t lcount_in_cap;
	double4* pos_prune;                     // [cap] positions at the last prune (displacement criterion of the inner lists)
	int      inner_valid;                   // the force kernels read lists_in / lcount_in; the guard also covers the displacement since the prune
	double   thr_in2;                       // (0.4995 skin_in)^2: the inner lists are exact while max |r - r_prune|^2 <= this
	uint64_t list_prunes;
	int      inner_life, since_prune;       // steps the inner lists lived when they last expired (0 = unknown); steps since the last prune
	int      inner_streak;                  // proactive prunes since the last expiry (the expected lifetime is raised again after 16)
	int      vi_every_step;                 // LS1G_VI_EVERY_STEP=1: per-molecule virial on every list step, not only the last of a run
	int      guard_on;                      // force / integrate kernels return at once when d_flags[3] > disp_thr2 (lists expired)
	double   disp_thr2;                     // (0.4995 skin)^2: the lists are exact while max |r - r_build|^2 <= this
	cudaEvent_t disp_ev[16];                // completion of the per-step copy of d_flags[0] into h_flags[8 + step % 16]
	unsigned long long* h_flags;            // pinned [48]: 0, 2 build flags, 1 overflow bits, 3 boundary blocks, 4 near-pair flag, 8..23 ring of the
	                                        // displacement flags (since the build), 24..39 the same since the last prune      // device staging area of the host<->device transfers (transfer.cu)

	// cells / sort scratch
	uint32_t *cell_count, *cell_fill, *cell_start;   // [n_cells+1]
	uint32_t *scan_block;                            // block sums of the scan
	uint32_t *slot_src;  uint8_t* slot_code; uint64_t* slot_key; uint32_t* slot_cell;   // [cap]
	uint32_t *d_counters;                            // [8] misc device counters
	uint32_t *h_counters;                            // pinned mirror

	// reductions
	double*  partials; size_t partials_cap;          // per-block partial sums
	double*  d_scalars;                              // [16] reduced results
	double*  h_scalars;                              // pinned mirror
	unsigned long long* d_pairs;                     // pair counter
	uint64_t last_pairs;

	// multi-GPU
	void* nccl_comm; int n_ranks, rank;
	int comm_mode;                          // 0 single rank, 1 NCCL, 2 in-process local group (tests)
	double *send_buf[2], *recv_buf[2]; size_t xbuf_cap;
	// persistent halo lists of the neighbour-list loop (comm.cu: comm_exchange_build / comm_refresh), per stage and direction
	uint32_t *xl_send[3][2], *xl_recv[3][2];   // device: slots to send / slots the received positions go to (post-sort numbering)
	size_t   xl_send_cap[3][2], xl_recv_cap[3][2];
	uint32_t xl_n_send[3][2], xl_n_recv[3][2], xl_recv_base[3][2];
	uint32_t *xl_pre[2];                       // device scratch: pre-sort slot of every packed record (current stage)
	// in-process transport (comm_mode 2): the ranks are contexts of one process, driven in lockstep by one host thread
	std::vector<ls1g_ctx*> peers;              // all ranks of the local group in rank order (this context included)
	cudaEvent_t ev_pack, ev_read;              // send buffers of the current stage packed / neighbours' buffers copied into my receive buffers
	cudaEvent_t ev_ar;                         // my contribution to the current group reduction has reached gbuf
	double* gbuf; int gbuf_owner;              // pinned host ring [4][n_ranks][8] of the group reductions (allocated by rank 0)
	unsigned gseq;                             // group reductions issued so far (ring slot = gseq % 4)
	// per-step records of ls1g_run / ls1g_step on the device (copied out once per call)
	double* d_rec; size_t d_rec_cap;

	Timer timers[LS1G_T_COUNT];
	std::vector<PendingEvt> pending;        // event pairs not yet resolved into timers[]
	std::vector<cudaEvent_t> evpool;
};

// One rank (single context, with or without NCCL) or all ranks of an in-process group, driven in lockstep
struct Group {
	ls1g_ctx* const* c; int n;
	bool multi() const { return c[0]->n_ranks > 1; }
};

// error handling --------------------------------------------------------------------------------
void ls1g_set_error(const std::string& msg);
#define CUDA_TRY(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { \
	ls1g_set_error(std::string(#expr) + ": " + cudaGetErrorString(_e)); return 1; } } while (0)
#define LS1G_FAIL(msg) do { ls1g_set_error(msg); return 1; } while (0)

// launchers implemented in the kernel files ---------------------------------------------------------
// cells.cu
int  grid_init_cut(Grid& g, const ls1g_config& cfg, double cut, int mode);   // api.cu
int  k_halo_map(ls1g_ctx* c);                   // after k_sort(c, 1): halo_src / halo_code of every halo slot
int  k_cell_index(ls1g_ctx* c, size_t n, const double* d_pts, uint32_t* d_out);
int  k_sort(ls1g_ctx* c, int make_images);      // bin (+periodic images) -> scan -> scatter -> gather, swaps cur/alt
int  k_drop_halo(ls1g_ctx* c);                  // deleteOuterParticles
int  k_debug_cell_pairs(const ls1g_ctx* c, uint32_t* h_triplets, size_t capacity, size_t* n_out);
int  ls1g_scan_cells(ls1g_ctx* c, const uint32_t* in, uint32_t* out);   // exclusive scan over n_cells entries -> out[n_cells+1]
// transfer.cu
int  k_upload(ls1g_ctx* c, size_t n, const uint64_t* id, const uint32_t* cid, const double* r, const double* v, const double* q, const double* L);
int  k_upload_forces(ls1g_ctx* c, size_t n, const double* F, const double* M);
int  k_download(ls1g_ctx* c, int mask, int include_halo, size_t n_ret, ls1g_fields* out);
int  ls1g_nl_leave(ls1g_ctx* c);               // step.cu: back to the reference-grid state (no-op outside list mode)
// integrate.cu
int  k_scale(ls1g_ctx* c, double bt, double br);
int  k_post_pre(ls1g_ctx* c, double dt, int do_post, int do_pre, int thermostat, double target_T);
// force.cu
int  k_forces(ls1g_ctx* c);                           // results in d_scalars[0..3]; caller syncs
// peak.cu
int  k_fp64_peak(int device, double* tflops);
// comm.cu
int  comm_exchange(const Group& G);            // leaving + halo exchange of every rank of the group, then the local sort
int  comm_exchange_build(const Group& G);      // list mode: migrate leavers, then halo pass with persistent lists, sort
int  comm_refresh(const Group& G);             // list mode: refresh halo positions along the persistent lists (asynchronous)
enum { COMM_SUM = 0, COMM_MAX = 2 };
enum { COMM_SCALARS = 0, COMM_FLAGS = 1 };
// guard of the speculatively enqueued list steps: kernels return at once when *ptr > thr (lists.cuh).  With inner lists the
// criterion is the displacement since the last prune, else since the build.
static inline unsigned long long ls1g_double_bits(double x) { unsigned long long b; memcpy(&b, &x, sizeof(b)); return b; }
static inline const unsigned long long* nl_guard_ptr(const ls1g_ctx* c) { return c->guard_on ? c->d_flags + 3 : nullptr; }
static inline unsigned long long nl_guard_thr(const ls1g_ctx*) { return ls1g_double_bits(1.0); }      // d_flags[3] holds 0.0 (go) or 2.0 (expired), lists.cuh: nl_guard_kernel
bool nl_inner_wanted(const ls1g_ctx* c);      // lists.cu
int  k_list_prune(ls1g_ctx* c, int guarded);   // lists.cu
// all-reduce of n <= 8 doubles at d_scalars + off (COMM_SCALARS) or d_flags + off (COMM_FLAGS) of every rank, in place, asynchronous
int  comm_allreduce(const Group& G, int which, int off, int n, int op);
void comm_destroy(ls1g_ctx* c);

void timer_start(ls1g_ctx* c, int which);
void timer_stop(ls1g_ctx* c, int which, int launches);

// ------------------------------------------------------------------------------------------------
// Device helpers shared by kernels
// ------------------------------------------------------------------------------------------------
#ifdef __CUDACC__

__device__ __forceinline__ uint32_t pos_cid(const double4& p) { return (uint32_t)__double_as_longlong(p.w); }
__device__ __forceinline__ double   cid_bits(uint32_t cid) { return __longlong_as_double((long long)cid); }

// CellBorderAndFlagManager::getBorder (CellBorderAndFlagManager.h:134-157).  The general branch is an
// FMA in the reference's -O3 -mfma build (pinned by tests/golden/cellindex_*.npz).
__device__ __host__ __forceinline__ double grid_border(const Grid& g, int index, int d) {
	if (index == 0) return g.halo_min[d];
	if (index == 1) return g.box_min[d];
	if (index == g.cpd[d] - 1) return g.box_max[d];
	if (index == g.cpd[d]) return g.halo_max[d];
#ifdef __CUDA_ARCH__
	return __fma_rn((double)index, g.cell_len[d], g.halo_min[d]);
#else
	return __builtin_fma((double)index, g.cell_len[d], g.halo_min[d]);
#endif
}

// LinkedCells::getCellIndexOfPoint (LinkedCells.cpp:830-882): bit-exact (explicit rounding intrinsics,
// no contraction).
__device__ __host__ __forceinline__ uint32_t grid_cell_index(const Grid& g, double px, double py, double pz) {
	double p[3] = {px, py, pz};
	int ci[3];
#pragma unroll
	for (int d = 0; d < 3; ++d) {
		if (p[d] <= g.halo_min[d]) p[d] = p[d] + g.cell_len[d] * 0.5;
		else if (p[d] >= g.halo_max[d]) p[d] = p[d] - g.cell_len[d] * 0.5;
#ifdef __CUDA_ARCH__
		double t = __dmul_rn(__dsub_rn(p[d], g.box_min[d]), g.cell_len_recip[d]);
#else
		volatile double t0 = p[d] - g.box_min[d];
		volatile double t = t0 * g.cell_len_recip[d];
#endif
		int c = (int)floor(t) + 1;
		c = c < 0 ? 0 : c;
		c = c > g.cpd[d] - 1 ? g.cpd[d] - 1 : c;
		ci[d] = c;
	}
	bool in = true;
#pragma unroll
	for (int d = 0; d < 3; ++d) in = in && (grid_border(g, ci[d], d) <= p[d]) && (p[d] < grid_border(g, ci[d] + 1, d));
	if (!in) {
#pragma unroll
		for (int d = 0; d < 3; ++d) {
			int c0 = ci[d];
			if (p[d] < grid_border(g, c0, d)) ci[d] = c0 - 1;
			else if (p[d] >= grid_border(g, c0 + 1, d)) ci[d] = c0 + 1;
		}
	}
	return (uint32_t)((ci[2] * g.cpd[1] + ci[1]) * g.cpd[0] + ci[0]);
}

__device__ __host__ __forceinline__ bool grid_is_halo(const Grid& g, uint32_t cell) {
	int x = cell % g.cpd[0];
	int y = (cell / g.cpd[0]) % g.cpd[1];
	int z = cell / (g.cpd[0] * g.cpd[1]);
	return x == 0 || y == 0 || z == 0 || x == g.cpd[0] - 1 || y == g.cpd[1] - 1 || z == g.cpd[2] - 1;
}

// Quaternion::rotate (src/molecules/Quaternion.cpp:43-62)
__device__ __forceinline__ void quat_rotate(const double4& q, const double d[3], double out[3]) {
	double ww = q.x*q.x, xx = q.y*q.y, yy = q.z*q.z, zz = q.w*q.w;
	double wx = q.x*q.y, wy = q.x*q.z, wz = q.x*q.w, xy = q.y*q.z, xz = q.y*q.w, yz = q.z*q.w;
	out[0] = (ww+xx-yy-zz)*d[0] + 2.*(xy-wz)*d[1] + 2.*(wy+xz)*d[2];
	out[1] = 2.*(wz+xy)*d[0] + (ww-xx+yy-zz)*d[1] + 2.*(yz-wx)*d[2];
	out[2] = 2.*(xz-wy)*d[0] + 2.*(wx+yz)*d[1] + (ww-xx-yy+zz)*d[2];
}
// Quaternion::rotateinv (Quaternion.cpp:64-83)
__device__ __forceinline__ void quat_rotateinv(const double4& q, const double d[3], double out[3]) {
	double ww = q.x*q.x, xx = q.y*q.y, yy = q.z*q.z, zz = q.w*q.w;
	double wx = q.x*q.y, wy = q.x*q.z, wz = q.x*q.w, xy = q.y*q.z, xz = q.y*q.w, yz = q.z*q.w;
	out[0] = (ww+xx-yy-zz)*d[0] + 2.*(xy+wz)*d[1] + 2.*(xz-wy)*d[2];
	out[1] = 2.*(xy-wz)*d[0] + (ww-xx+yy-zz)*d[1] + 2.*(yz+wx)*d[2];
	out[2] = 2.*(xz+wy)*d[0] + 2.*(yz-wx)*d[1] + (ww-xx-yy+zz)*d[2];
}

// warp / block sum of `NV` doubles; result valid in thread 0 of the block.  Fixed order -> deterministic.
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* smem /* [NV*32] */) {
#pragma unroll
	for (int k = 0; k < NV; ++k)
#pragma unroll
		for (int o = 16; o > 0; o >>= 1) v[k] += __shfl_down_sync(0xffffffffu, v[k], o);
	int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
	if (lane == 0)
#pragma unroll
		for (int k = 0; k < NV; ++k) smem[k * 32 + warp] = v[k];
	__syncthreads();
	if (warp == 0) {
#pragma unroll
		for (int k = 0; k < NV; ++k) {
			double x = lane < nw ? smem[k * 32 + lane] : 0.0;
#pragma unroll
			for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
			v[k] = x;
		}
	}
	__syncthreads();
}
#endif  // __CUDACC__
