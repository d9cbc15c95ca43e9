// records.cu — the reference's binary phase-space records on the device boundary (ls1g_upload_records / ls1g_download_records).
//
// Replaces (reference, src/): the record loop of BinaryReader::readPhaseSpace (io/BinaryReader.cpp:171-287: formats ICRVQD / ICRV /
// IRV :183-217, component id check :222-228, Molecule constructor + isInBoundingBox + addParticle :232-273, maxid :281-282) on the
// way in, and DomainDecompBase::writeMoleculesToFile -> FullMolecule::writeBinary (parallel/DomainDecompBase.cpp:505-540,
// molecules/FullMolecule.cpp:451-474) on the way out.
//
// A record is PACKED: id (8 B) | component id, 1-based (4 B) | r v (6 x 8 B) | q = w x y z (4 x 8 B) | D (3 x 8 B) = 116 bytes, so
// the doubles sit on 4-byte boundaries.  The host never touches a record: the byte stream of the file is copied into the device
// staging area chunk by chunk as it is, a block moves 256 records (29 696 contiguous bytes) between HBM and shared memory with
// 16-byte accesses, and every thread (de)composes one record from 32-bit words in shared memory (record stride 29 words: odd, no
// bank conflicts).  HBM-bound byte work: 116 B read + (32 pos + 24 vel + 8 id + 4 cell [+ 32 q + 24 D]) written per molecule.
#include "ls1g_internal.cuh"

static inline unsigned nblk(size_t n, unsigned t) { return (unsigned)((n + t - 1) / t); }

#define REC_BLOCK 256                       // records per block
#define REC_CHUNK (REC_BLOCK * 1024)        // records per chunk of the staging area: one 1024-thread block scans its block counts
static const int REC_WORDS[3] = {29, 15, 14};   // LS1G_REC_ICRVQD, LS1G_REC_ICRV, LS1G_REC_IRV in 32-bit words

// control block behind the record bytes of the staging area
struct RecCtl {
	unsigned long long kept;      // records kept so far (all chunks before the current one)
	unsigned long long max_id;    // BinaryReader.cpp:281: highest id of ALL records, kept or not
	unsigned long long bad;       // records with a component id of 0 or beyond the component count
	unsigned long long overflow;  // kept records that did not fit the capacity
	uint32_t blk_off[REC_CHUNK / REC_BLOCK + 1];   // kept records before each block of the current chunk
};

struct RecBox { double lo[3], hi[3]; };

// global -> shared: n_words 32-bit words from a 16-byte aligned source
__device__ __forceinline__ void rec_load_words(uint32_t* sh, const uint32_t* __restrict__ src, int n_words) {
	const int nq = n_words >> 2;
	const uint4* s4 = reinterpret_cast<const uint4*>(src);
	uint4* d4 = reinterpret_cast<uint4*>(sh);
	for (int i = threadIdx.x; i < nq; i += REC_BLOCK) d4[i] = __ldg(s4 + i);
	for (int i = (nq << 2) + threadIdx.x; i < n_words; i += REC_BLOCK) sh[i] = __ldg(src + i);
}

__device__ __forceinline__ double rec_f64(const uint32_t* w) { return __hiloint2double((int)w[1], (int)w[0]); }

// ParticleContainer::isInBoundingBox (particleContainer/ParticleContainer.cpp:37-44)
__device__ __forceinline__ bool rec_in_box(const RecBox& b, double x, double y, double z) {
	return x >= b.lo[0] && y >= b.lo[1] && z >= b.lo[2] && x < b.hi[0] && y < b.hi[1] && z < b.hi[2];
}

// offset of r inside a record, in words: behind id (+ component id)
__device__ __forceinline__ int rec_r_word(int W) { return W == 14 ? 2 : 3; }

// pass 1 over a chunk: records of every block that lie inside the bounding box
__global__ void __launch_bounds__(REC_BLOCK) rec_count_kernel(int n, int W, const uint32_t* __restrict__ rec, RecBox box, RecCtl* __restrict__ ctl) {
	__shared__ int warp_cnt[REC_BLOCK / 32];
	const int i = blockIdx.x * REC_BLOCK + threadIdx.x;
	bool keep = false;
	if (i < n) {
		const uint32_t* w = rec + (size_t)i * W + rec_r_word(W);
		uint32_t v[6];
#pragma unroll
		for (int k = 0; k < 6; ++k) v[k] = __ldg(w + k);
		keep = rec_in_box(box, rec_f64(v), rec_f64(v + 2), rec_f64(v + 4));
	}
	const unsigned m = __ballot_sync(0xffffffffu, keep);
	if ((threadIdx.x & 31) == 0) warp_cnt[threadIdx.x >> 5] = __popc(m);
	__syncthreads();
	if (threadIdx.x == 0) {
		int s = 0;
		for (int k = 0; k < REC_BLOCK / 32; ++k) s += warp_cnt[k];
		ctl->blk_off[blockIdx.x] = (uint32_t)s;
	}
}

// exclusive scan of the block counts of a chunk (one block of 1024 threads), in place; blk_off[n_blocks] = kept records of the chunk
__global__ void __launch_bounds__(1024) rec_scan_kernel(int n_blocks, RecCtl* __restrict__ ctl) {
	__shared__ uint32_t warp_sum[32];
	const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
	const uint32_t v = t < n_blocks ? ctl->blk_off[t] : 0u;
	uint32_t x = v;
#pragma unroll
	for (int d = 1; d < 32; d <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
	if (lane == 31) warp_sum[wid] = x;
	__syncthreads();
	if (wid == 0) {
		uint32_t s = warp_sum[lane];
#pragma unroll
		for (int d = 1; d < 32; d <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, s, d); if (lane >= d) s += y; }
		warp_sum[lane] = s;
	}
	__syncthreads();
	const uint32_t incl = x + (wid ? warp_sum[wid - 1] : 0u);
	if (t < n_blocks) ctl->blk_off[t] = incl - v;
	if (t == n_blocks - 1) ctl->blk_off[n_blocks] = incl;
}

// the chunk is done: its kept records join the running total
__global__ void rec_advance_kernel(int n_blocks, RecCtl* __restrict__ ctl) { ctl->kept += ctl->blk_off[n_blocks]; }

// pass 2: records -> molecule arrays, kept records in file order behind those of the earlier chunks
__global__ void __launch_bounds__(REC_BLOCK) rec_unpack_kernel(int n, int W, const uint32_t* __restrict__ rec, RecBox box, RecCtl* __restrict__ ctl,
                                                               MolBuf mol, int has_rot, uint32_t n_comp, size_t cap) {
	__shared__ __align__(16) uint32_t sh[REC_BLOCK * 29];
	__shared__ int warp_cnt[REC_BLOCK / 32];
	const int first = blockIdx.x * REC_BLOCK;
	const int cnt = min(REC_BLOCK, n - first);
	rec_load_words(sh, rec + (size_t)first * W, cnt * W);
	__syncthreads();
	const int t = threadIdx.x, lane = t & 31, wid = t >> 5;
	const uint32_t* w = sh + t * W;
	bool keep = false;
	unsigned long long id = 0ull;
	uint32_t cid = 1u;
	double x = 0., y = 0., z = 0.;
	if (t < cnt) {
		id = (unsigned long long)w[0] | ((unsigned long long)w[1] << 32);
		const int o = rec_r_word(W);
		if (W != 14) cid = w[2];                       // IRV: componentid = 1 (BinaryReader.cpp:207)
		x = rec_f64(w + o); y = rec_f64(w + o + 2); z = rec_f64(w + o + 4);
		keep = rec_in_box(box, x, y, z);
	}
	// every record counts for maxid and for the component id check, kept or not (BinaryReader.cpp:222-228,281-282)
	unsigned long long mx = id;
#pragma unroll
	for (int d = 16; d; d >>= 1) { unsigned long long o = __shfl_xor_sync(0xffffffffu, mx, d); mx = o > mx ? o : mx; }
	const bool bad = t < cnt && (cid == 0u || cid > n_comp);
	const unsigned mbad = __ballot_sync(0xffffffffu, bad);
	const unsigned mk = __ballot_sync(0xffffffffu, keep);
	if (lane == 0) {
		warp_cnt[wid] = __popc(mk);
		if (mx) atomicMax(&ctl->max_id, mx);
		if (mbad) atomicAdd(&ctl->bad, (unsigned long long)__popc(mbad));
	}
	__syncthreads();
	if (!keep) return;
	int rank = __popc(mk & ((1u << lane) - 1u));
	for (int k = 0; k < wid; ++k) rank += warp_cnt[k];
	const size_t dst = (size_t)ctl->kept + ctl->blk_off[blockIdx.x] + (size_t)rank;
	if (dst >= cap) { atomicAdd(&ctl->overflow, 1ull); return; }
	const int o = rec_r_word(W);
	mol.pos[dst] = make_double4(x, y, z, cid_bits(bad ? 0u : cid - 1u));      // component ids start with 0 in the program (BinaryReader.cpp:229)
	mol.vel[3 * dst] = rec_f64(w + o + 6); mol.vel[3 * dst + 1] = rec_f64(w + o + 8); mol.vel[3 * dst + 2] = rec_f64(w + o + 10);
	mol.id[dst] = id;
	mol.cell[dst] = 0u;
	if (has_rot) {
		if (W == 29) {
			mol.q[dst] = make_double4(rec_f64(w + 15), rec_f64(w + 17), rec_f64(w + 19), rec_f64(w + 21));     // (w x y z) in the double4's (x y z w) slots
			mol.L[3 * dst] = rec_f64(w + 23); mol.L[3 * dst + 1] = rec_f64(w + 25); mol.L[3 * dst + 2] = rec_f64(w + 27);
		} else {                                       // q0 = 1, q1 = q2 = q3 = Dx = Dy = Dz = 0 (BinaryReader.cpp:167-168)
			mol.q[dst] = make_double4(1., 0., 0., 0.);
			mol.L[3 * dst] = 0.; mol.L[3 * dst + 1] = 0.; mol.L[3 * dst + 2] = 0.;
		}
	}
}

__global__ void rec_ctl_clear_kernel(RecCtl* __restrict__ ctl) { ctl->kept = 0ull; ctl->max_id = 0ull; ctl->bad = 0ull; ctl->overflow = 0ull; }

// staging area: room for one chunk of records and the control block behind it (grow-only, shared with transfer.cu)
static int rec_stage(ls1g_ctx* c, size_t n, int words, size_t* chunk, RecCtl** ctl) {
	const size_t per = (size_t)words * 4;
	size_t ch = std::min<size_t>(std::max<size_t>(n, 1), REC_CHUNK);
	const size_t ctl_off = (ch * per + 255) & ~(size_t)255;
	const size_t want = std::max(ctl_off + sizeof(RecCtl), c->cap * sizeof(double4));
	if (!c->stage || c->stage_bytes < want) {
		cudaFree(c->stage); c->stage = nullptr; c->stage_bytes = 0;
		CUDA_TRY(cudaMalloc(&c->stage, want));
		c->stage_bytes = want;
	}
	*chunk = ch;
	*ctl = reinterpret_cast<RecCtl*>((char*)c->stage + ctl_off);
	return 0;
}

int k_upload_records(ls1g_ctx* c, size_t n, const void* records, int format, uint64_t* n_kept, uint64_t* max_id) {
	const int W = REC_WORDS[format];
	size_t chunk; RecCtl* ctl;
	if (rec_stage(c, n, W, &chunk, &ctl)) return 1;
	cudaStream_t st = c->stream;
	RecBox box;
	for (int d = 0; d < 3; ++d) { box.lo[d] = c->cfg.box_min[d]; box.hi[d] = c->cfg.box_max[d]; }
	rec_ctl_clear_kernel<<<1, 1, 0, st>>>(ctl);
	for (size_t k0 = 0; k0 < n; k0 += chunk) {
		const int m = (int)std::min(chunk, n - k0);
		const int nb = (int)nblk(m, REC_BLOCK);
		CUDA_TRY(cudaMemcpyAsync(c->stage, (const char*)records + k0 * W * 4, (size_t)m * W * 4, cudaMemcpyHostToDevice, st));
		rec_count_kernel<<<nb, REC_BLOCK, 0, st>>>(m, W, (const uint32_t*)c->stage, box, ctl);
		rec_scan_kernel<<<1, 1024, 0, st>>>(nb, ctl);
		rec_unpack_kernel<<<nb, REC_BLOCK, 0, st>>>(m, W, (const uint32_t*)c->stage, box, ctl, c->cur, c->has_rot, (uint32_t)c->n_comp, c->cap);
		rec_advance_kernel<<<1, 1, 0, st>>>(nb, ctl);
	}
	unsigned long long h[4] = {0, 0, 0, 0};
	CUDA_TRY(cudaMemcpyAsync(h, ctl, sizeof(h), cudaMemcpyDeviceToHost, st));
	CUDA_TRY(cudaStreamSynchronize(st));
	CUDA_TRY(cudaGetLastError());
	if (h[2]) LS1G_FAIL("ls1g_upload_records: " + std::to_string(h[2]) + " record(s) with a wrong component id (0 or above the component count)");
	if (h[3]) LS1G_FAIL("ls1g_upload_records: " + std::to_string(h[0]) + " molecules inside the bounding box, capacity " + std::to_string(c->cap));
	const size_t kept = (size_t)h[0];
	if (kept) {
		CUDA_TRY(cudaMemsetAsync(c->F, 0, kept * 3 * sizeof(double), st));
		if (c->has_rot) CUDA_TRY(cudaMemsetAsync(c->M, 0, kept * 3 * sizeof(double), st));
		CUDA_TRY(cudaStreamSynchronize(st));
	}
	*n_kept = h[0];
	*max_id = h[1];
	return 0;
}

// --------------------------------------------------------------------------------------------------
// owned molecules of every cell: 0 for halo cells (the rows of ls1g_download without halo copies)
__global__ void __launch_bounds__(256) rec_kept_kernel(Grid g, const uint32_t* __restrict__ cell_start, uint32_t* __restrict__ out) {
	const int cidx = blockIdx.x * blockDim.x + threadIdx.x;
	if (cidx >= g.n_cells) return;
	out[cidx] = grid_is_halo(g, (uint32_t)cidx) ? 0u : cell_start[cidx + 1] - cell_start[cidx];
}

// rows [row0, row0 + n) of the owned molecules in cell order -> ICRVQD records.  kept_before == nullptr: row == slot.
__global__ void __launch_bounds__(REC_BLOCK) rec_pack_kernel(size_t row0, int n, MolBuf mol, int has_rot, double sv, double sd, const uint32_t* __restrict__ cell_start,
                                                             const uint32_t* __restrict__ kept_before, int n_cells, uint32_t* __restrict__ rec) {
	__shared__ __align__(16) uint32_t sh[REC_BLOCK * 29];
	const int first = blockIdx.x * REC_BLOCK;
	const int cnt = min(REC_BLOCK, n - first);
	const int t = threadIdx.x;
	if (t < cnt) {
		const size_t row = row0 + first + t;
		size_t s = row;
		if (kept_before) {
			// the cell of this row: the last one with kept_before[cell] <= row (cells that keep nothing share the value of their successor)
			int lo = 0, hi = n_cells;                 // first index in (lo, hi] with kept_before[index] > row
			while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (__ldg(kept_before + mid) > (uint32_t)row) hi = mid; else lo = mid; }
			s = (size_t)__ldg(cell_start + lo) + (row - __ldg(kept_before + lo));
		}
		uint32_t* w = sh + t * 29;
		auto put = [&](int k, double v) { w[k] = (uint32_t)__double2loint(v); w[k + 1] = (uint32_t)__double2hiint(v); };
		const unsigned long long id = mol.id[s];
		const double4 p = mol.pos[s];
		w[0] = (uint32_t)id; w[1] = (uint32_t)(id >> 32);
		w[2] = pos_cid(p) + 1u;                       // FullMolecule::writeBinary: _component->ID() + 1
		put(3, p.x); put(5, p.y); put(7, p.z);
		put(9, mol.vel[3 * s] * sv); put(11, mol.vel[3 * s + 1] * sv); put(13, mol.vel[3 * s + 2] * sv);      // Molecule::scale_v: v[d] *= s (1.0 is exact)
		if (has_rot) {
			const double4 q = mol.q[s];
			put(15, q.x); put(17, q.y); put(19, q.z); put(21, q.w);
			put(23, mol.L[3 * s] * sd); put(25, mol.L[3 * s + 1] * sd); put(27, mol.L[3 * s + 2] * sd);           // Molecule::scale_D
		} else {
			put(15, 1.); put(17, 0.); put(19, 0.); put(21, 0.); put(23, 0.); put(25, 0.); put(27, 0.);
		}
	}
	__syncthreads();
	const int n_words = cnt * 29, nq = n_words >> 2;
	uint32_t* dst = rec + (size_t)first * 29;           // 256 records = 29 696 bytes: every block starts on a 16-byte boundary
	uint4* d4 = reinterpret_cast<uint4*>(dst);
	const uint4* s4 = reinterpret_cast<const uint4*>(sh);
	for (int i = t; i < nq; i += REC_BLOCK) d4[i] = s4[i];
	for (int i = (nq << 2) + t; i < n_words; i += REC_BLOCK) dst[i] = sh[i];
}

int k_download_records(ls1g_ctx* c, double scale_v, double scale_D, size_t n_ret, void* records) {
	if (n_ret == 0) return 0;
	size_t chunk; RecCtl* ctl;
	if (rec_stage(c, n_ret, 29, &chunk, &ctl)) return 1;
	cudaStream_t st = c->stream;
	const uint32_t* kept = nullptr;
	if (n_ret != c->n_slots) {                          // halo copies in between: rows skip the halo cells (as ls1g_download does)
		rec_kept_kernel<<<nblk(c->bgrid.n_cells, 256), 256, 0, st>>>(c->bgrid, c->cell_start, c->cell_count);
		if (ls1g_scan_cells(c, c->cell_count, c->cell_fill)) return 1;
		kept = c->cell_fill;
	}
	for (size_t k0 = 0; k0 < n_ret; k0 += chunk) {
		const int m = (int)std::min(chunk, n_ret - k0);
		rec_pack_kernel<<<nblk(m, REC_BLOCK), REC_BLOCK, 0, st>>>(k0, m, c->cur, c->has_rot, scale_v, scale_D, c->cell_start, kept, c->bgrid.n_cells, (uint32_t*)c->stage);
		CUDA_TRY(cudaMemcpyAsync((char*)records + k0 * 116, c->stage, (size_t)m * 116, cudaMemcpyDeviceToHost, st));
	}
	CUDA_TRY(cudaStreamSynchronize(st));
	CUDA_TRY(cudaGetLastError());
	return 0;
}
