// comm.cu — block domain decomposition across the GPUs of one box: leaving + halo exchange and scalar sums.
//
// Replaces (reference, src/parallel/):
//   DomainDecomposition::balanceAndExchange            DomainDecomposition.cpp:87-103
//   IndirectNeighbourCommunicationScheme (x, y, z)      NeighbourCommunicationScheme.cpp:500-600 ("indirect": corners via faces)
//   CommunicationPartner::initSend / collectMoleculesInRegion   CommunicationPartner.cpp:139-222,430-560
//   CommunicationBuffer record packing                  CommunicationBuffer.cpp:27-36,131-176
//   Domain::calculateGlobalValues collComm allreduce    Domain.cpp:170-176,214-224
//
// Semantics kept from the reference (full-shell, "leaving together with halo copies"):
//   stage d = 0,1,2; for both directions: every particle present whose coordinate d lies OUTSIDE the box
//   towards that neighbour is sent (it stays here as a halo particle: removeFromContainer=false,
//   CommunicationPartner.h:55), every particle inside the rc-wide boundary slab is sent as a halo copy.
//   Particles received in stage d take part in the later stages, so edges/corners travel through faces.
//   The sender applies the periodic shift +-L when the neighbour is across the global boundary, with the
//   reference's rounding guards (CommunicationPartner.cpp:530-560).  Ownership is geometric: after the
//   last stage the counting sort puts every particle into its cell; halo cells = halo particles.
// Dimensions with one rank (self-periodic) are handled by the fused image path of the sort (cells.cu).
//
// Transport: NCCL send/recv over NVLink (one process per GPU), loaded at run time from the torch-bundled
// libnccl.so.2; or an in-process "local" transport (several contexts in one process, device-to-device
// copies) used to test the exchange logic on a single GPU.
#include "ls1g_internal.cuh"

#include <dlfcn.h>
#include <string.h>
#include <stdlib.h>
#include <vector>
#include <algorithm>

// ---- minimal NCCL surface (signatures from nccl.h 2.2x; loaded with dlsym) ------------------------
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef int ncclResult_t;
enum { NCCL_DOUBLE = 8, NCCL_UINT64 = 5, NCCL_SUM = 0 };
struct NcclApi {
	void* lib;
	ncclResult_t (*GetUniqueId)(ncclUniqueId*);
	ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int);
	ncclResult_t (*CommDestroy)(ncclComm_t);
	ncclResult_t (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t);
	ncclResult_t (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t);
	ncclResult_t (*AllReduce)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t);
	ncclResult_t (*GroupStart)();
	ncclResult_t (*GroupEnd)();
	const char* (*GetErrorString)(ncclResult_t);
};
static NcclApi g_nccl = {nullptr};

static int nccl_load() {
	if (g_nccl.lib) return 0;
	const char* env = getenv("LS1G_NCCL_LIB");
	const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
	void* h = nullptr;
	for (const char* n : names) { if (!n) continue; h = dlopen(n, RTLD_NOW | RTLD_GLOBAL); if (h) break; }
	if (!h) LS1G_FAIL(std::string("ls1g: cannot load NCCL (set LS1G_NCCL_LIB to libnccl.so.2): ") + dlerror());
#define SYM(field, name) do { *(void**)(&g_nccl.field) = dlsym(h, name); if (!g_nccl.field) LS1G_FAIL(std::string("ls1g: NCCL symbol missing: ") + name); } while (0)
	SYM(GetUniqueId, "ncclGetUniqueId"); SYM(CommInitRank, "ncclCommInitRank"); SYM(CommDestroy, "ncclCommDestroy");
	SYM(Send, "ncclSend"); SYM(Recv, "ncclRecv"); SYM(AllReduce, "ncclAllReduce");
	SYM(GroupStart, "ncclGroupStart"); SYM(GroupEnd, "ncclGroupEnd"); SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
	g_nccl.lib = h;
	return 0;
}
#define NCCL_TRY(expr) do { ncclResult_t _r = (expr); if (_r != 0) LS1G_FAIL(std::string(#expr) + ": " + g_nccl.GetErrorString(_r)); } while (0)

// ------------------------------------------------------------------------------------------------
// record = 8 doubles (x y z cid | vx vy vz id), +8 with rotation (qw qx qy qz | Lx Ly Lz 0); when the forces of the sender are
// valid (an exchange outside the time loop, e.g. on leaving the neighbour-list layout: the next half-kick needs F(t) of a molecule
// that changes its rank) +4 (Fx Fy Fz 0) and, with rotation, +4 (Mx My Mz 0)
// ------------------------------------------------------------------------------------------------
struct PackArgs {
	Grid g;
	size_t n;
	MolBuf m;
	int has_rot, dim, rec;    // rec = doubles per record
	const double *F, *M;      // carried when non-null
	double shift[2];          // periodic shift applied for direction 0 (lower) / 1 (upper); 0 = none
	double global_len;
	double* out[2]; uint32_t* counters; uint32_t cap;
	int mode;                 // 0: leaving + halo in one pass (the reference's scheme); 1: leaving only, sent molecules are removed
	                          // here; 2: halo copies only, the source slot of every record is remembered in rec_slot
	uint32_t* rec_slot[2];
};

__device__ __forceinline__ void write_record(const PackArgs& a, int dir, size_t i, double4 p) {
	uint32_t k = atomicAdd(&a.counters[dir], 1u);
	if (k >= a.cap) return;                     // overflow: detected on the host from the counter
	if (a.mode == 2) a.rec_slot[dir][k] = (uint32_t)i;
	double* o = a.out[dir] + (size_t)k * a.rec;
	o[0] = p.x; o[1] = p.y; o[2] = p.z; o[3] = p.w;
	o[4] = a.m.vel[3 * i]; o[5] = a.m.vel[3 * i + 1]; o[6] = a.m.vel[3 * i + 2];
	o[7] = __longlong_as_double((long long)a.m.id[i]);
	if (a.has_rot) {
		double4 q = a.m.q[i];
		o[8] = q.x; o[9] = q.y; o[10] = q.z; o[11] = q.w;
		o[12] = a.m.L[3 * i]; o[13] = a.m.L[3 * i + 1]; o[14] = a.m.L[3 * i + 2]; o[15] = 0.;
	}
	if (a.F) {
		double* f = o + (a.has_rot ? 16 : 8);
		f[0] = a.F[3 * i]; f[1] = a.F[3 * i + 1]; f[2] = a.F[3 * i + 2]; f[3] = 0.;
		if (a.M) { f[4] = a.M[3 * i]; f[5] = a.M[3 * i + 1]; f[6] = a.M[3 * i + 2]; f[7] = 0.; }
	}
}

__global__ void __launch_bounds__(256) pack_kernel(PackArgs a) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= a.n) return;
	if (a.m.cell[i] & HALO_FLAG) return;          // stale halo copy of the previous step
	const Grid& g = a.g;
	const int d = a.dim;
	double4 p = a.m.pos[i];
	double x = d == 0 ? p.x : (d == 1 ? p.y : p.z);
	for (int dir = 0; dir < 2; ++dir) {
		bool leaving = dir == 0 ? (x < g.box_min[d]) : (x >= g.box_max[d]);
		double lo = dir == 0 ? g.box_min[d] : __dsub_rn(g.box_max[d], g.cutoff);
		double hi = dir == 0 ? __dadd_rn(g.box_min[d], g.cutoff) : g.box_max[d];
		bool slab = x >= lo && x < hi;
		if (a.mode == 1) slab = false;
		if (a.mode == 2) leaving = false;
		if (!(leaving || slab)) continue;
		double y = x;
		double sh = a.shift[dir];
		if (sh != 0.) {
			y = __dadd_rn(x, sh);
			if (slab) {            // halo copy must land outside the global box (CommunicationPartner.cpp:533-548)
				if (sh < 0.) { if (y >= 0.) y = nextafter(0., -1.); }
				else { if (y < a.global_len) y = nextafter(a.global_len, a.global_len + 1.); }
			} else {               // leaving particle must land inside (:549-560)
				if (sh < 0.) { if (y < 0.) y = 0.; }
				else { if (y >= a.global_len) y = nextafter(a.global_len, a.global_len - 1.); }
			}
		}
		double4 o = p;
		if (d == 0) o.x = y; else if (d == 1) o.y = y; else o.z = y;
		write_record(a, dir, i, o);
		if (a.mode == 1) a.m.cell[i] = HALO_FLAG;      // migrated: the next sort drops this slot
	}
}

// ---- persistent halo lists (neighbour-list loop) --------------------------------------------------
__global__ void __launch_bounds__(256) translate_slots_kernel(size_t n, const uint32_t* __restrict__ pre, uint32_t base, const uint32_t* __restrict__ new_of_old,
                                                               uint32_t* __restrict__ post) {
	size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (k < n) post[k] = new_of_old[pre ? pre[k] : base + (uint32_t)k];
}
// positions only: 3 doubles per record (7 with the orientation of a rigid molecule), shifted along the stage dimension when the
// neighbour is across the global boundary
__global__ void __launch_bounds__(256) pack_pos_list_kernel(size_t n, const uint32_t* __restrict__ slots, const double4* __restrict__ pos, const double4* __restrict__ q,
                                                             int dim, double shift, double* __restrict__ out) {
	size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	const uint32_t s = slots[k];
	double4 p = pos[s];
	if (dim == 0) p.x += shift; else if (dim == 1) p.y += shift; else p.z += shift;
	const int w = q ? 7 : 3;
	out[w * k] = p.x; out[w * k + 1] = p.y; out[w * k + 2] = p.z;
	if (q) { const double4 qq = q[s]; out[w * k + 3] = qq.x; out[w * k + 4] = qq.y; out[w * k + 5] = qq.z; out[w * k + 6] = qq.w; }
}
__global__ void __launch_bounds__(256) unpack_pos_list_kernel(size_t n, const uint32_t* __restrict__ slots, const double* __restrict__ in, double4* __restrict__ pos,
                                                               double4* __restrict__ q) {
	size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	const uint32_t s = slots[k];
	const int w = q ? 7 : 3;
	double4 p = pos[s];
	p.x = in[w * k]; p.y = in[w * k + 1]; p.z = in[w * k + 2];
	pos[s] = p;
	if (q) q[s] = make_double4(in[w * k + 3], in[w * k + 4], in[w * k + 5], in[w * k + 6]);
}

__global__ void __launch_bounds__(256) unpack_kernel(const double* __restrict__ in, size_t n, size_t base, MolBuf m, int has_rot, int rec,
                                                      double* __restrict__ F, double* __restrict__ M) {
	size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (k >= n) return;
	const double* r = in + k * rec;
	size_t i = base + k;
	m.pos[i] = make_double4(r[0], r[1], r[2], r[3]);
	m.vel[3 * i] = r[4]; m.vel[3 * i + 1] = r[5]; m.vel[3 * i + 2] = r[6];
	m.id[i] = (uint64_t)__double_as_longlong(r[7]);
	m.cell[i] = 0;
	if (has_rot) {
		m.q[i] = make_double4(r[8], r[9], r[10], r[11]);
		m.L[3 * i] = r[12]; m.L[3 * i + 1] = r[13]; m.L[3 * i + 2] = r[14];
	}
	if (F) {
		const double* f = r + (has_rot ? 16 : 8);
		F[3 * i] = f[0]; F[3 * i + 1] = f[1]; F[3 * i + 2] = f[2];
		if (M) { M[3 * i] = f[4]; M[3 * i + 1] = f[5]; M[3 * i + 2] = f[6]; }
	}
}

__global__ void __launch_bounds__(256) count_owned_kernel(size_t n, const uint32_t* __restrict__ cell, uint32_t* counter) {
	size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
	bool owned = i < n && !(cell[i] & HALO_FLAG);
	unsigned b = __ballot_sync(0xffffffffu, owned);
	if ((threadIdx.x & 31) == 0 && b) atomicAdd(counter, (uint32_t)__popc(b));
}

// ------------------------------------------------------------------------------------------------
static inline unsigned nblk(size_t n, unsigned t) { return (unsigned)((n + t - 1) / t); }
// forces travel with the molecules whenever they are valid (k_sort carries them through the re-sort under the same condition)
static inline int rec_doubles(const ls1g_ctx* c) { return (c->has_rot ? 16 : 8) + (c->forces_valid ? (c->has_rot ? 8 : 4) : 0); }

static int rank_of(const int grid[3], int cx, int cy, int cz) { return (cx * grid[1] + cy) * grid[2] + cz; }   // MPI_Cart_rank, row-major

static int neighbour_rank(const ls1g_ctx* c, int d, int dir) {
	int co[3] = {c->cfg.rank_coords[0], c->cfg.rank_coords[1], c->cfg.rank_coords[2]};
	const int* g = c->cfg.rank_grid;
	co[d] = (co[d] + (dir == 0 ? -1 : 1) + g[d]) % g[d];
	return rank_of(g, co[0], co[1], co[2]);
}

static int ensure_xbuf(ls1g_ctx* c) {
	// records per direction and stage: everything a small system holds may have to travel; of a large one only the boundary slab
	// does (a stage that needs more fails with "send buffer overflow")
	size_t want = c->cap < (1u << 20) ? c->cap : std::max<size_t>(c->cap / 4, 1u << 20);
	if (c->xbuf_cap >= want && c->send_buf[0]) return 0;
	for (int k = 0; k < 2; ++k) { cudaFree(c->send_buf[k]); cudaFree(c->recv_buf[k]); c->send_buf[k] = c->recv_buf[k] = nullptr; }
	size_t bytes = want * (c->has_rot ? 24 : 12) * sizeof(double);
	for (int k = 0; k < 2; ++k) {
		CUDA_TRY(cudaMalloc((void**)&c->send_buf[k], bytes));
		CUDA_TRY(cudaMalloc((void**)&c->recv_buf[k], bytes));
		cudaFree(c->xl_pre[k]); c->xl_pre[k] = nullptr;
		CUDA_TRY(cudaMalloc((void**)&c->xl_pre[k], want * sizeof(uint32_t)));
	}
	c->xbuf_cap = want;
	return 0;
}

// phase 1 of stage d: pack both directions; send counts end in h_counters[2], [3]
static int stage_pack(ls1g_ctx* c, int d, int mode = 0) {
	if (ensure_xbuf(c)) return 1;
	CUDA_TRY(cudaMemsetAsync(c->d_counters + 2, 0, 2 * sizeof(uint32_t), c->stream));
	PackArgs a;
	a.g = c->bgrid; a.n = c->n_slots; a.m = c->cur; a.has_rot = c->has_rot; a.dim = d; a.rec = rec_doubles(c);
	a.F = c->forces_valid ? c->F : nullptr; a.M = c->forces_valid && c->has_rot ? c->M : nullptr;
	a.global_len = c->cfg.global_length[d];
	a.shift[0] = c->cfg.rank_coords[d] == 0 ? c->cfg.global_length[d] : 0.;
	a.shift[1] = c->cfg.rank_coords[d] == c->cfg.rank_grid[d] - 1 ? -c->cfg.global_length[d] : 0.;
	a.out[0] = c->send_buf[0]; a.out[1] = c->send_buf[1]; a.counters = c->d_counters + 2; a.cap = (uint32_t)c->xbuf_cap;
	a.mode = mode; a.rec_slot[0] = c->xl_pre[0]; a.rec_slot[1] = c->xl_pre[1];
	if (c->n_slots) pack_kernel<<<nblk(c->n_slots, 256), 256, 0, c->stream>>>(a);
	CUDA_TRY(cudaMemcpyAsync(c->h_counters + 2, c->d_counters + 2, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
	CUDA_TRY(cudaStreamSynchronize(c->stream));
	if (c->h_counters[2] > c->xbuf_cap || c->h_counters[3] > c->xbuf_cap) LS1G_FAIL("ls1g exchange: send buffer overflow (raise ls1g_config.capacity)");
	return 0;
}

// phase 3: append received records
static int stage_unpack(ls1g_ctx* c, const uint32_t n_recv[2]) {
	for (int k = 0; k < 2; ++k) {
		if (!n_recv[k]) continue;
		if (c->n_slots + n_recv[k] > c->cap) LS1G_FAIL("ls1g exchange: capacity exceeded by received particles (raise ls1g_config.capacity)");
		unpack_kernel<<<nblk(n_recv[k], 256), 256, 0, c->stream>>>(c->recv_buf[k], n_recv[k], c->n_slots, c->cur, c->has_rot, rec_doubles(c),
		                                                         c->forces_valid ? c->F : nullptr, c->forces_valid && c->has_rot ? c->M : nullptr);
		c->n_slots += n_recv[k];
	}
	CUDA_TRY(cudaGetLastError());
	return 0;
}

static int finish_exchange(ls1g_ctx* c) {
	// self-periodic dimensions + final cell sort
	if (k_sort(c, 1)) return 1;
	CUDA_TRY(cudaMemsetAsync(c->d_counters + 1, 0, sizeof(uint32_t), c->stream));
	if (c->n_slots) count_owned_kernel<<<nblk(c->n_slots, 256), 256, 0, c->stream>>>(c->n_slots, c->cur.cell, c->d_counters + 1);
	CUDA_TRY(cudaMemcpyAsync(c->h_counters + 1, c->d_counters + 1, sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
	CUDA_TRY(cudaStreamSynchronize(c->stream));
	c->n_owned = c->h_counters[1];
	c->n_halo = c->n_slots - c->n_owned;
	return 0;
}

static int check_grid_order(const ls1g_ctx* c) {
	bool seen_self = false;
	for (int d = 0; d < 3; ++d) {
		if (c->cfg.rank_grid[d] <= 1) seen_self = true;
		else if (seen_self) LS1G_FAIL("ls1g: rank grid must be non-increasing (decomposed dimensions first), as ls1g_decomp_grid returns it");
	}
	return 0;
}

// ---- transports ------------------------------------------------------------------------------------
// NCCL: one process per GPU, the group is this rank alone and the neighbours are reached by ncclSend/ncclRecv.
// local: all ranks are contexts of this process (any number of GPUs >= 1); records move by device-to-device copies, ordered by
//        events.  Everything around the transfer itself - pack / unpack kernels, persistent lists, slot translation, the
//        overlapped two-phase force launch, the guarded steps - is the same code for both.
static inline void set_dev(const ls1g_ctx* c) { cudaSetDevice(c->cfg.device); }

// one stage (dimension d) of a NCCL rank: pack, exchange counts and records with both neighbours, append the received records.
// n_recv_out[2]: records received from the upper / lower neighbour; *base_out: slot of the first appended record.
static int nccl_stage(ls1g_ctx* c, int d, int mode, uint32_t n_recv_out[2], uint32_t* base_out) {
	ncclComm_t comm = (ncclComm_t)c->nccl_comm;
	const int rec = rec_doubles(c);
	if (stage_pack(c, d, mode)) return 1;
	const int lower = neighbour_rank(c, d, 0), upper = neighbour_rank(c, d, 1);
	// counts: d_counters[2],[3] = my sends (lower, upper); [4] = from upper, [5] = from lower
	NCCL_TRY(g_nccl.GroupStart());
	NCCL_TRY(g_nccl.Send(c->d_counters + 2, 1, 3 /*ncclUint32*/, lower, comm, c->stream));
	NCCL_TRY(g_nccl.Send(c->d_counters + 3, 1, 3, upper, comm, c->stream));
	NCCL_TRY(g_nccl.Recv(c->d_counters + 4, 1, 3, upper, comm, c->stream));
	NCCL_TRY(g_nccl.Recv(c->d_counters + 5, 1, 3, lower, comm, c->stream));
	NCCL_TRY(g_nccl.GroupEnd());
	CUDA_TRY(cudaMemcpyAsync(c->h_counters + 4, c->d_counters + 4, 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
	CUDA_TRY(cudaStreamSynchronize(c->stream));
	uint32_t n_recv[2] = {c->h_counters[4], c->h_counters[5]};
	if (n_recv[0] > c->xbuf_cap || n_recv[1] > c->xbuf_cap) LS1G_FAIL("ls1g exchange: receive buffer overflow (raise ls1g_config.capacity)");
	NCCL_TRY(g_nccl.GroupStart());
	if (c->h_counters[2]) NCCL_TRY(g_nccl.Send(c->send_buf[0], (size_t)c->h_counters[2] * rec, NCCL_DOUBLE, lower, comm, c->stream));
	if (c->h_counters[3]) NCCL_TRY(g_nccl.Send(c->send_buf[1], (size_t)c->h_counters[3] * rec, NCCL_DOUBLE, upper, comm, c->stream));
	if (n_recv[0]) NCCL_TRY(g_nccl.Recv(c->recv_buf[0], (size_t)n_recv[0] * rec, NCCL_DOUBLE, upper, comm, c->stream));
	if (n_recv[1]) NCCL_TRY(g_nccl.Recv(c->recv_buf[1], (size_t)n_recv[1] * rec, NCCL_DOUBLE, lower, comm, c->stream));
	NCCL_TRY(g_nccl.GroupEnd());
	*base_out = (uint32_t)c->n_slots;
	if (stage_unpack(c, n_recv)) return 1;
	n_recv_out[0] = n_recv[0]; n_recv_out[1] = n_recv[1];
	return 0;
}

static int check_group(const Group& G) {
	ls1g_ctx* c0 = G.c[0];
	if (c0->n_ranks <= 1) return 0;
	if (c0->comm_mode == 1) { if (G.n != 1) LS1G_FAIL("ls1g: a NCCL rank is driven alone (one process per GPU)"); return 0; }
	if (c0->comm_mode != 2) LS1G_FAIL("ls1g: multi-rank context without a transport (ls1g_comm_init / ls1g_comm_init_local)");
	if (G.n != c0->n_ranks || (int)c0->peers.size() != G.n) LS1G_FAIL("ls1g: a local group is driven as a whole (all ranks, in rank order)");
	for (int r = 0; r < G.n; ++r) if (G.c[r] != c0->peers[r]) LS1G_FAIL("ls1g: contexts do not match the group of ls1g_comm_init_local");
	return 0;
}

// one stage of every rank of the group (blocking).  nr[r][2], base[r] as nccl_stage.
struct NR2 { uint32_t v[2]; uint32_t& operator[](int k) { return v[k]; } };
static int group_stage(const Group& G, int d, int mode, NR2* nr, uint32_t* base) {
	if (G.c[0]->comm_mode == 1) return nccl_stage(G.c[0], d, mode, nr[0].v, &base[0]);
	for (int r = 0; r < G.n; ++r) { set_dev(G.c[r]); if (stage_pack(G.c[r], d, mode)) return 1; }      // synchronises: counts in h_counters[2], [3]
	for (int r = 0; r < G.n; ++r) {
		ls1g_ctx* c = G.c[r];
		set_dev(c);
		const int rec = rec_doubles(c);
		ls1g_ctx* up = G.c[neighbour_rank(c, d, 1)];
		ls1g_ctx* lo = G.c[neighbour_rank(c, d, 0)];
		// from my upper neighbour I get what it sent downwards (its buffer 0); from my lower neighbour its buffer 1
		nr[r][0] = up->h_counters[2]; nr[r][1] = lo->h_counters[3];
		if (nr[r][0] > c->xbuf_cap || nr[r][1] > c->xbuf_cap) LS1G_FAIL("ls1g exchange: receive buffer overflow (raise ls1g_config.capacity)");
		if (nr[r][0]) CUDA_TRY(cudaMemcpyAsync(c->recv_buf[0], up->send_buf[0], (size_t)nr[r][0] * rec * sizeof(double), cudaMemcpyDefault, c->stream));
		if (nr[r][1]) CUDA_TRY(cudaMemcpyAsync(c->recv_buf[1], lo->send_buf[1], (size_t)nr[r][1] * rec * sizeof(double), cudaMemcpyDefault, c->stream));
	}
	for (int r = 0; r < G.n; ++r) {
		ls1g_ctx* c = G.c[r];
		set_dev(c);
		base[r] = (uint32_t)c->n_slots;
		if (stage_unpack(c, nr[r].v)) return 1;
		CUDA_TRY(cudaStreamSynchronize(c->stream));   // send buffers are reused by the next stage
	}
	return 0;
}

int comm_exchange(const Group& G) {
	if (check_group(G)) return 1;
	if (check_grid_order(G.c[0])) return 1;
	std::vector<uint32_t> base(G.n);
	std::vector<NR2> nr(G.n);
	for (int r = 0; r < G.n; ++r) { set_dev(G.c[r]); timer_start(G.c[r], LS1G_T_EXCHANGE); }
	int launches = 0;
	for (int d = 0; d < 3; ++d) {
		if (G.c[0]->cfg.rank_grid[d] <= 1) continue;
		if (group_stage(G, d, 0, nr.data(), base.data())) return 1;
		launches += 4;
	}
	for (int r = 0; r < G.n; ++r) { set_dev(G.c[r]); timer_stop(G.c[r], LS1G_T_EXCHANGE, launches); }
	for (int r = 0; r < G.n; ++r) { set_dev(G.c[r]); if (finish_exchange(G.c[r])) return 1; }
	return 0;
}

static int ensure_u32(uint32_t** p, size_t* cap, size_t want) {
	if (*cap >= want && *p) return 0;
	cudaFree(*p); *p = nullptr; *cap = 0;
	CUDA_TRY(cudaMalloc((void**)p, (want + want / 8 + 1024) * sizeof(uint32_t)));
	*cap = want + want / 8 + 1024;
	return 0;
}

// Exchange of a list-build step.  The reference sends leaving molecules together with the halo copies and keeps the
// leaver's old copy as the sender's halo particle (CommunicationPartner.h:55); with a frozen layout that copy would never be
// refreshed, so here ownership moves first (pass 1), then the halo copies are created (pass 2) and every stage remembers
// which slots it sent and where the received copies live: comm_refresh re-sends just their positions on later steps.
int comm_exchange_build(const Group& G) {
	if (check_group(G)) return 1;
	if (check_grid_order(G.c[0])) return 1;
	std::vector<uint32_t> base(G.n);
	std::vector<NR2> nr(G.n);
	for (int r = 0; r < G.n; ++r) { set_dev(G.c[r]); timer_start(G.c[r], LS1G_T_EXCHANGE); }
	int launches = 0;
	for (int d = 0; d < 3; ++d) {
		if (G.c[0]->cfg.rank_grid[d] <= 1) continue;
		if (group_stage(G, d, 1, nr.data(), base.data())) return 1;
		launches += 4;
	}
	// pre-sort bookkeeping of pass 2, translated after the sort
	for (int d = 0; d < 3; ++d) {
		for (int r = 0; r < G.n; ++r)
			for (int k = 0; k < 2; ++k) { G.c[r]->xl_n_send[d][k] = G.c[r]->xl_n_recv[d][k] = 0; G.c[r]->xl_recv_base[d][k] = 0; }
		if (G.c[0]->cfg.rank_grid[d] <= 1) continue;
		if (group_stage(G, d, 2, nr.data(), base.data())) return 1;
		launches += 4;
		for (int r = 0; r < G.n; ++r) {
			ls1g_ctx* c = G.c[r];
			set_dev(c);
			for (int k = 0; k < 2; ++k) {
				c->xl_n_send[d][k] = c->h_counters[2 + k];
				if (ensure_u32(&c->xl_send[d][k], &c->xl_send_cap[d][k], c->xl_n_send[d][k])) return 1;
				// keep the pre-sort slots in xl_send until they are translated
				if (c->xl_n_send[d][k]) CUDA_TRY(cudaMemcpyAsync(c->xl_send[d][k], c->xl_pre[k], c->xl_n_send[d][k] * sizeof(uint32_t), cudaMemcpyDeviceToDevice, c->stream));
			}
			c->xl_n_recv[d][0] = nr[r][0]; c->xl_n_recv[d][1] = nr[r][1];
			c->xl_recv_base[d][0] = base[r]; c->xl_recv_base[d][1] = base[r] + nr[r][0];
			for (int k = 0; k < 2; ++k) if (ensure_u32(&c->xl_recv[d][k], &c->xl_recv_cap[d][k], c->xl_n_recv[d][k])) return 1;
			CUDA_TRY(cudaStreamSynchronize(c->stream));      // xl_pre is rewritten by the next stage's pack
		}
	}
	for (int r = 0; r < G.n; ++r) {
		ls1g_ctx* c = G.c[r];
		set_dev(c);
		timer_stop(c, LS1G_T_EXCHANGE, launches);
		if (finish_exchange(c)) return 1;
		const uint32_t* new_of_old = c->alt.cell;     // valid directly after k_sort (cells.cu: gather_kernel)
		for (int d = 0; d < 3; ++d)
			for (int k = 0; k < 2; ++k) {
				if (c->xl_n_send[d][k])
					translate_slots_kernel<<<nblk(c->xl_n_send[d][k], 256), 256, 0, c->stream>>>(c->xl_n_send[d][k], c->xl_send[d][k], 0, new_of_old, c->xl_send[d][k]);
				if (c->xl_n_recv[d][k])
					translate_slots_kernel<<<nblk(c->xl_n_recv[d][k], 256), 256, 0, c->stream>>>(c->xl_n_recv[d][k], nullptr, c->xl_recv_base[d][k], new_of_old, c->xl_recv[d][k]);
			}
		CUDA_TRY(cudaGetLastError());
	}
	return 0;
}

// frozen-layout step: positions of the halo copies follow their owners, stage by stage (x, y, z: forwarded copies are
// refreshed before they are forwarded again).  Asynchronous on every rank's current stream (no host synchronisation).
int comm_refresh(const Group& G) {
	const bool local = G.c[0]->comm_mode == 2;
	for (int r = 0; r < G.n; ++r) { set_dev(G.c[r]); timer_start(G.c[r], LS1G_T_EXCHANGE); }
	std::vector<int> launches(G.n, 0);
	for (int d = 0; d < 3; ++d) {
		if (G.c[0]->cfg.rank_grid[d] <= 1) continue;
		for (int r = 0; r < G.n; ++r) {
			ls1g_ctx* c = G.c[r];
			set_dev(c);
			const double shift[2] = {c->cfg.rank_coords[d] == 0 ? c->cfg.global_length[d] : 0.,
			                         c->cfg.rank_coords[d] == c->cfg.rank_grid[d] - 1 ? -c->cfg.global_length[d] : 0.};
			if (local) {      // my send buffers are free once both neighbours have copied the previous stage out of them
				CUDA_TRY(cudaStreamWaitEvent(c->stream, G.c[neighbour_rank(c, d, 1)]->ev_read, 0));
				CUDA_TRY(cudaStreamWaitEvent(c->stream, G.c[neighbour_rank(c, d, 0)]->ev_read, 0));
			}
			for (int k = 0; k < 2; ++k)
				if (c->xl_n_send[d][k]) {
					pack_pos_list_kernel<<<nblk(c->xl_n_send[d][k], 256), 256, 0, c->stream>>>(c->xl_n_send[d][k], c->xl_send[d][k], c->cur.pos, c->has_rot ? c->cur.q : nullptr, d, shift[k], c->send_buf[k]);
					++launches[r];
				}
			if (local) CUDA_TRY(cudaEventRecord(c->ev_pack, c->stream));
		}
		for (int r = 0; r < G.n; ++r) {
			ls1g_ctx* c = G.c[r];
			set_dev(c);
			const int lower = neighbour_rank(c, d, 0), upper = neighbour_rank(c, d, 1);
			const size_t pw = c->has_rot ? 7 : 3;      // doubles per refreshed halo molecule
			if (local) {
				ls1g_ctx *up = G.c[upper], *lo = G.c[lower];
				CUDA_TRY(cudaStreamWaitEvent(c->stream, up->ev_pack, 0));
				CUDA_TRY(cudaStreamWaitEvent(c->stream, lo->ev_pack, 0));
				if (c->xl_n_recv[d][0]) CUDA_TRY(cudaMemcpyAsync(c->recv_buf[0], up->send_buf[0], (size_t)c->xl_n_recv[d][0] * pw * sizeof(double), cudaMemcpyDefault, c->stream));
				if (c->xl_n_recv[d][1]) CUDA_TRY(cudaMemcpyAsync(c->recv_buf[1], lo->send_buf[1], (size_t)c->xl_n_recv[d][1] * pw * sizeof(double), cudaMemcpyDefault, c->stream));
				CUDA_TRY(cudaEventRecord(c->ev_read, c->stream));
			} else {
				ncclComm_t comm = (ncclComm_t)c->nccl_comm;
				NCCL_TRY(g_nccl.GroupStart());
				if (c->xl_n_send[d][0]) NCCL_TRY(g_nccl.Send(c->send_buf[0], (size_t)c->xl_n_send[d][0] * pw, NCCL_DOUBLE, lower, comm, c->stream));
				if (c->xl_n_send[d][1]) NCCL_TRY(g_nccl.Send(c->send_buf[1], (size_t)c->xl_n_send[d][1] * pw, NCCL_DOUBLE, upper, comm, c->stream));
				if (c->xl_n_recv[d][0]) NCCL_TRY(g_nccl.Recv(c->recv_buf[0], (size_t)c->xl_n_recv[d][0] * pw, NCCL_DOUBLE, upper, comm, c->stream));
				if (c->xl_n_recv[d][1]) NCCL_TRY(g_nccl.Recv(c->recv_buf[1], (size_t)c->xl_n_recv[d][1] * pw, NCCL_DOUBLE, lower, comm, c->stream));
				NCCL_TRY(g_nccl.GroupEnd());
			}
			for (int k = 0; k < 2; ++k)
				if (c->xl_n_recv[d][k]) {
					unpack_pos_list_kernel<<<nblk(c->xl_n_recv[d][k], 256), 256, 0, c->stream>>>(c->xl_n_recv[d][k], c->xl_recv[d][k], c->recv_buf[k], c->cur.pos, c->has_rot ? c->cur.q : nullptr);
					++launches[r];
				}
		}
	}
	for (int r = 0; r < G.n; ++r) { set_dev(G.c[r]); CUDA_TRY(cudaGetLastError()); timer_stop(G.c[r], LS1G_T_EXCHANGE, launches[r]); }
	return 0;
}

// out[k] = op over ranks of g[r][k]; g is the pinned host ring slot every rank copied its values into (zero-copy reads)
__global__ void group_reduce_kernel(const double* g, int n_ranks, int n, int op, double* __restrict__ out) {
	const int k = threadIdx.x;
	if (k >= n) return;
	double v = __ldcv(g + k);
	for (int r = 1; r < n_ranks; ++r) { const double x = __ldcv(g + 8 * r + k); v = op == COMM_MAX ? fmax(v, x) : v + x; }
	out[k] = v;
}

int comm_allreduce(const Group& G, int which, int off, int n, int op) {
	ls1g_ctx* c0 = G.c[0];
	if (c0->n_ranks <= 1) return 0;
	if (n < 1 || n > 8) LS1G_FAIL("ls1g: group reduction of more than 8 values");
	auto ptr = [&](ls1g_ctx* c) { return which == COMM_FLAGS ? (double*)c->d_flags + off : c->d_scalars + off; };
	if (c0->comm_mode == 1) {
		set_dev(c0);
		NCCL_TRY(g_nccl.AllReduce(ptr(c0), ptr(c0), (size_t)n, NCCL_DOUBLE, op == COMM_MAX ? 2 /* ncclMax */ : NCCL_SUM, (ncclComm_t)c0->nccl_comm, c0->stream));
		return 0;
	}
	if (c0->comm_mode != 2 || !c0->gbuf) LS1G_FAIL("ls1g: group reduction without a transport");
	// ring of 4 slots: a rank can only be a full reduction ahead of the slowest one (every reduction waits for everybody's
	// contribution), so slot s is not rewritten while somebody still reads it
	double* slot = c0->gbuf + (size_t)(c0->gseq % 4) * G.n * 8;
	for (int r = 0; r < G.n; ++r) {
		ls1g_ctx* c = G.c[r];
		set_dev(c);
		CUDA_TRY(cudaMemcpyAsync(slot + 8 * r, ptr(c), n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
		CUDA_TRY(cudaEventRecord(c->ev_ar, c->stream));
	}
	for (int r = 0; r < G.n; ++r) {
		ls1g_ctx* c = G.c[r];
		set_dev(c);
		for (int q = 0; q < G.n; ++q) if (q != r) CUDA_TRY(cudaStreamWaitEvent(c->stream, G.c[q]->ev_ar, 0));
		group_reduce_kernel<<<1, 32, 0, c->stream>>>(slot, G.n, n, op, ptr(c));
		CUDA_TRY(cudaGetLastError());
	}
	for (int r = 0; r < G.n; ++r) G.c[r]->gseq = c0->gseq + 1;
	return 0;
}

extern "C" int ls1g_comm_unique_id(unsigned char id[128]) {
	if (nccl_load()) return 1;
	ncclUniqueId u;
	NCCL_TRY(g_nccl.GetUniqueId(&u));
	memcpy(id, u.internal, 128);
	return 0;
}

extern "C" int ls1g_comm_init(ls1g_ctx* c, int n_ranks, int rank, const unsigned char id[128]) {
	if (!c) LS1G_FAIL("null context");
	if (n_ranks != c->cfg.rank_grid[0] * c->cfg.rank_grid[1] * c->cfg.rank_grid[2]) LS1G_FAIL("ls1g_comm_init: n_ranks does not match rank_grid");
	if (rank != rank_of(c->cfg.rank_grid, c->cfg.rank_coords[0], c->cfg.rank_coords[1], c->cfg.rank_coords[2]))
		LS1G_FAIL("ls1g_comm_init: rank does not match rank_coords (row-major, as ls1g_decomp_grid)");
	if (check_grid_order(c)) return 1;
	c->n_ranks = n_ranks; c->rank = rank;
	if (n_ranks == 1) { c->comm_mode = 0; return 0; }
	if (nccl_load()) return 1;
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	ncclUniqueId u; memcpy(u.internal, id, 128);
	ncclComm_t comm;
	NCCL_TRY(g_nccl.CommInitRank(&comm, n_ranks, u, rank));
	c->nccl_comm = comm; c->comm_mode = 1;
	return 0;
}

void comm_destroy(ls1g_ctx* c) {
	if (c->comm_mode == 1 && c->nccl_comm && g_nccl.lib) g_nccl.CommDestroy((ncclComm_t)c->nccl_comm);
	if (c->comm_mode == 2) {
		// the group dissolves: the surviving members fall back to "no transport" (their next collective call fails loudly)
		for (ls1g_ctx* p : c->peers) if (p != c) { p->peers.clear(); if (p->gbuf && !p->gbuf_owner) p->gbuf = nullptr; p->comm_mode = 0; }
		if (c->gbuf && c->gbuf_owner) cudaFreeHost(c->gbuf);
		else for (ls1g_ctx* p : c->peers) if (p != c && p->gbuf_owner && p->gbuf) { cudaFreeHost(p->gbuf); p->gbuf = nullptr; p->gbuf_owner = 0; }
		c->gbuf = nullptr; c->gbuf_owner = 0; c->peers.clear();
	}
	c->nccl_comm = nullptr; c->comm_mode = 0;
}

extern "C" int ls1g_allreduce_sum(ls1g_ctx* c, double* values, int n) {
	if (!c || !values || n < 0 || n > 8) LS1G_FAIL("ls1g_allreduce_sum: bad arguments (n <= 8)");
	if (c->n_ranks <= 1) return 0;
	if (c->comm_mode != 1) LS1G_FAIL("ls1g_allreduce_sum: NCCL ranks only (a local group lives in one process: sum on the host)");
	CUDA_TRY(cudaSetDevice(c->cfg.device));
	double* d = c->d_scalars + 8;
	CUDA_TRY(cudaMemcpyAsync(d, values, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
	Group G{&c, 1};
	if (comm_allreduce(G, COMM_SCALARS, 8, n, COMM_SUM)) return 1;
	CUDA_TRY(cudaMemcpyAsync(values, d, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
	CUDA_TRY(cudaStreamSynchronize(c->stream));
	return 0;
}

// ---- local transport: all ranks are contexts of this process (any number of GPUs >= 1) ----------------------
extern "C" int ls1g_comm_init_local(ls1g_ctx** ctxs, int n) {
	if (!ctxs || n < 1) LS1G_FAIL("ls1g_comm_init_local: bad arguments");
	for (int r = 0; r < n; ++r) {
		ls1g_ctx* c = ctxs[r];
		if (!c) LS1G_FAIL("null context");
		if (n != c->cfg.rank_grid[0] * c->cfg.rank_grid[1] * c->cfg.rank_grid[2]) LS1G_FAIL("ls1g_comm_init_local: n does not match rank_grid");
		if (r != rank_of(c->cfg.rank_grid, c->cfg.rank_coords[0], c->cfg.rank_coords[1], c->cfg.rank_coords[2]))
			LS1G_FAIL("ls1g_comm_init_local: contexts must be ordered by rank (row-major coords)");
		if (check_grid_order(c)) return 1;
		if (c->comm_mode != 0) LS1G_FAIL("ls1g_comm_init_local: context already has a transport");
	}
	double* gbuf = nullptr;
	if (n > 1) {
		CUDA_TRY(cudaSetDevice(ctxs[0]->cfg.device));
		CUDA_TRY(cudaHostAlloc((void**)&gbuf, (size_t)4 * n * 8 * sizeof(double), cudaHostAllocPortable | cudaHostAllocMapped));
		memset(gbuf, 0, (size_t)4 * n * 8 * sizeof(double));
		// contexts on different GPUs: direct copies between them (falls back to staged copies where peer access is unavailable)
		for (int r = 0; r < n; ++r)
			for (int q = 0; q < n; ++q) {
				const int a = ctxs[r]->cfg.device, b = ctxs[q]->cfg.device;
				if (a == b) continue;
				int can = 0;
				if (cudaDeviceCanAccessPeer(&can, a, b) == cudaSuccess && can) {
					cudaSetDevice(a);
					cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
					if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) LS1G_FAIL(std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e));
					cudaGetLastError();
				}
			}
	}
	for (int r = 0; r < n; ++r) {
		ls1g_ctx* c = ctxs[r];
		c->n_ranks = n; c->rank = r; c->comm_mode = n > 1 ? 2 : 0;
		c->peers.assign(ctxs, ctxs + n);
		c->gbuf = gbuf; c->gbuf_owner = r == 0 && gbuf; c->gseq = 0;
	}
	return 0;
}

// one leaving + halo exchange of all local ranks in lockstep (= ls1g_exchange on every rank)
int ls1g_nl_leave_group(const Group& G);
extern "C" int ls1g_exchange_group(ls1g_ctx** ctxs, int n) {
	if (!ctxs || n < 1) LS1G_FAIL("ls1g_exchange_group: bad arguments");
	if (n == 1) return ls1g_exchange(ctxs[0]);
	Group G{ctxs, n};
	if (check_group(G)) return 1;
	if (ls1g_nl_leave_group(G)) return 1;
	if (comm_exchange(G)) return 1;
	for (int r = 0; r < n; ++r) { set_dev(ctxs[r]); CUDA_TRY(cudaStreamSynchronize(ctxs[r]->stream)); }
	return 0;
}
