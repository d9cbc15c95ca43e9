/*
 * ls1g.h — C ABI of the B200 (sm_100a) short-range force + integrate path of ls1-MarDyn.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.  The C++ shim
 * classes a maintainer adds to ls1 (GpuLinkedCells : ParticleContainer, GpuLeapfrog : Integrator;
 * see INTEGRATION.md) forward to exactly these entry points.  Every entry point names the
 * reference interface it replaces (paths relative to the reference tree).
 *
 * Conventions
 *   - All pointers are caller-owned HOST memory; arrays of vectors are AoS ([n][3], [n][4]).
 *   - All calls are synchronous on return and return 0 on success, non-zero on error;
 *     ls1g_last_error() gives the message (the shim maps it to global_log->error + Simulation::exit,
 *     the reference's own error path: src/particleContainer/LinkedCells.cpp:101,162).
 *   - One context = one ParticleContainer = one GPU = one rank (src/Simulation.cpp:135-139).
 *   - There is no CPU fallback: without a CUDA device every compute call fails.
 */
#ifndef LS1G_H
#define LS1G_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ls1g_ctx ls1g_ctx;

/* How the cell grid is derived from the cutoff (the reference has two code paths). */
enum {
	LS1G_GRID_REBUILD = 0, /* LinkedCells::rebuild, XML path: float rc   (LinkedCells.cpp:136-204) */
	LS1G_GRID_CTOR    = 1  /* LinkedCells(min,max,rc), test path: double rc (LinkedCells.cpp:42-107) */
};

typedef struct {
	int    device;            /* CUDA device ordinal */
	double box_min[3];        /* bounding box of THIS rank: rebuild(bBoxMin,bBoxMax) (LinkedCells.cpp:136) */
	double box_max[3];
	double global_length[3];  /* Domain::getGlobalLength (periodic shift magnitude) */
	double cutoff;            /* rc   : charge/dipole/quadrupole + cell size (Simulation.cpp:257-267) */
	double lj_cutoff;         /* rcLJ : LJ sites */
	double eps_rf;            /* Domain::getepsilonRF (reaction field) */
	int    grid_mode;         /* LS1G_GRID_* */
	int    rank_grid[3];      /* DomainDecomposition::_gridSize  (DomainDecomposition.cpp:35) ; {1,1,1} = sequential */
	int    rank_coords[3];    /* DomainDecomposition::_coords */
	uint64_t capacity;        /* max molecules incl. halo copies this context may hold; 0 = grow on upload */
} ls1g_config;

/* One Component (src/molecules/Component.h).  Site tables are row-major doubles:
 *   lj         [n_lj][7]         x y z m eps sigma shift6     (Site.h:84-141)
 *   charge     [n_charge][5]     x y z m q                    (Site.h:143-193)
 *   dipole     [n_dipole][7]     x y z ex ey ez absMy         (Site.h:195-300)
 *   quadrupole [n_quadrupole][7] x y z ex ey ez absQ
 */
typedef struct {
	double mass;              /* Component::m()  */
	double inertia[3];        /* I11, I22, I33 (principal axes) */
	int    rot_dof;           /* Component::getRotationalDegreesOfFreedom() */
	int    n_lj, n_charge, n_dipole, n_quadrupole;
	const double* lj;
	const double* charge;
	const double* dipole;
	const double* quadrupole;
} ls1g_component;

/* Fields for ls1g_download (bit mask). */
enum {
	LS1G_F_ID = 1, LS1G_F_CID = 2, LS1G_F_R = 4, LS1G_F_V = 8, LS1G_F_Q = 16, LS1G_F_L = 32,
	LS1G_F_F = 64, LS1G_F_M = 128, LS1G_F_VI = 256, LS1G_F_CELL = 512
};

typedef struct {
	uint64_t* id;     /* [n]    */
	uint32_t* cid;    /* [n]    */
	double*   r;      /* [n][3] */
	double*   v;      /* [n][3] */
	double*   q;      /* [n][4] (w,x,y,z) */
	double*   L;      /* [n][3] angular momentum D */
	double*   F;      /* [n][3] */
	double*   M;      /* [n][3] */
	double*   Vi;     /* [n][3] */
	uint32_t* cell;   /* [n] linear cell index (z*ny+y)*nx+x incl. halo layer */
} ls1g_fields;

/* Per-step record of ls1g_run (one row per time step). */
typedef struct {
	double   upot;      /* Domain::setLocalUpot   value (VectorizedCellProcessor.cpp:156) */
	double   virial;    /* Domain::setLocalVirial value (VectorizedCellProcessor.cpp:155) */
	double   summv2;    /* Domain::setLocalSummv2 (Leapfrog.cpp:148) */
	double   sumIw2;    /* Domain::setLocalSumIw2 */
	uint64_t n;         /* Domain::setLocalNrotDOF N      */
	uint64_t rot_dof;   /* Domain::setLocalNrotDOF rotDOF */
} ls1g_step_record;

const char* ls1g_last_error(void);
int ls1g_version(void);
/* CUDA devices visible to this process (a container that spreads its box over several GPUs maps rank r to device r mod count).
 * replaces: the rank count of MPI_Comm_size (DomainDecompMPIBase.cpp:36-40) for a single-process multi-GPU run */
int ls1g_device_count(int* n);

/* --- life cycle --------------------------------------------------------------------------
 * replaces: Simulation::readXML datastructure dispatch + container->rebuild (Simulation.cpp:411-450) */
int ls1g_create(const ls1g_config* cfg, ls1g_ctx** out);
int ls1g_destroy(ls1g_ctx* ctx);

/* replaces: VectorizedCellProcessor ctor tables (VectorizedCellProcessor.cpp:21-98),
 * Comp2Param::initialize (Comp2Param.cpp:10-187), Ensemble::setComponentLookUpIDs (EnsembleBase.cpp:91-103).
 * mixcoeff = Domain::getmixcoeff(): (xi, eta) per component pair i<j, row-major upper triangle. */
int ls1g_set_components(ls1g_ctx* ctx, int n_components, const ls1g_component* comps,
                        const double* mixcoeff, int n_mixcoeff);

/* --- grid queries (host only, no GPU work) ------------------------------------------------
 * replaces: LinkedCells::rebuild geometry (LinkedCells.cpp:136-204), CellBorderAndFlagManager */
int ls1g_get_grid(const ls1g_ctx* ctx, int cells_per_dim[3], double cell_length[3]);
/* replaces: LinkedCells::getCellIndexOfPoint (LinkedCells.cpp:830-882); host evaluation */
int ls1g_cell_index_of_points_host(const ls1g_ctx* ctx, size_t n, const double* points, uint32_t* out);
/* same on the device (the kernel the sort uses) */
int ls1g_cell_index_of_points(ls1g_ctx* ctx, size_t n, const double* points, uint32_t* out);
/* NO counterpart in the reference: added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernel and same result as
 * ls1g_cell_index_of_points, but the call allocates and frees NOTHING: the device staging of `points` and of the result is a fixed
 * file-scope __device__ array (1.75 MiB, held by every GPU that loads the library until the process exits).
 * n > LS1G_NOALLOC_MAX_POINTS is rejected with an error.  The array is shared by every call and every context of a device and keeps
 * the last call's contents (nothing in it needs clearing: each call overwrites and reads back exactly [0, n)); calls that overlap on
 * the GPU race on it, so callers serialise. */
#define LS1G_NOALLOC_MAX_POINTS 65536
int ls1g_cell_index_of_points_noalloc(ls1g_ctx* ctx, size_t n, const double* points, uint32_t* out);
/* replaces: C08CellPairTraversal::traverseCellPairs + C08BasedTraversals::processBaseCell
 * (C08CellPairTraversal.h:50-80, C08BasedTraversals.h:48-102): emits (cell1, cell2, kind) triplets,
 * kind 0 = processCell, 1 = processCellPair, in colour/base-cell order.  pairs may be NULL to count. */
int ls1g_debug_cell_pairs(const ls1g_ctx* ctx, uint32_t* triplets, size_t capacity, size_t* n_out);

/* --- particle storage ---------------------------------------------------------------------
 * replaces: ParticleContainer::addParticles / clear (LinkedCells.cpp:445-520,596-601). Replaces the content. */
/* std::vector::reserve for the device: allocates the molecule arrays and the transfer staging for n owned molecules (plus the
 * halo copies the grid implies) ahead of ls1g_upload, which otherwise allocates on its first call.  Set-up cost, like ls1g_create. */
int ls1g_reserve(ls1g_ctx* ctx, size_t n);
int ls1g_upload(ls1g_ctx* ctx, size_t n, const uint64_t* id, const uint32_t* cid,
                const double* r, const double* v, const double* q, const double* L);
/* replaces: Molecule::setF / setM (molecules/FullMolecule.h) for a restart WITH forces: F (and M, may be NULL)
 * of the molecules of the directly preceding ls1g_upload, same order.  Makes ls1g_run callable without a
 * fresh ls1g_traverse (the first half-kick of Leapfrog::transition1to2 needs F(t), Leapfrog.cpp:48-68). */
int ls1g_upload_forces(ls1g_ctx* ctx, size_t n, const double* F, const double* M);
/* --- binary phase-space records ---------------------------------------------
 * The PACKED records of the reference's binary checkpoint, taken and produced by the device as they stand in the file: the
 * host copies bytes and never looks inside a record (no per-molecule host loop, no host Molecule objects).
 *   LS1G_REC_ICRVQD  id u64 | component id u32, 1-based | r[3] v[3] q[4] = (w x y z) D[3]   = 116 bytes
 *   LS1G_REC_ICRV    id u64 | component id u32, 1-based | r[3] v[3]                         =  60 bytes  (q = 1 0 0 0, D = 0)
 *   LS1G_REC_IRV     id u64 | r[3] v[3]                                                      =  56 bytes  (component 1)
 * replaces: <format type="ICRVQD|ICRV|IRV"/> of BinaryReader::readPhaseSpaceHeader (io/BinaryReader.cpp:102-112) */
enum { LS1G_REC_ICRVQD = 0, LS1G_REC_ICRV = 1, LS1G_REC_IRV = 2 };
int ls1g_record_size(int format, size_t* bytes);
/* replaces: the record loop of BinaryReader::readPhaseSpace (io/BinaryReader.cpp:171-287).  Replaces the content of the context by
 * the records that lie inside its bounding box (ParticleContainer::isInBoundingBox, ParticleContainer.cpp:37-44: min <= r < max, as
 * the reader's "if (isInBoundingBox) addParticle" :271-273 - every rank of a decomposed run is handed the whole file and keeps its
 * own part), in file order.  *n_kept = their number, *max_id = the highest id of ALL n records (the reader's return value, :281-282);
 * both may be NULL.  A component id of 0 or above the component count is an error (:222-228) and leaves the context empty.
 * With every record inside the box the state equals that after ls1g_upload of the same molecules (ls1g_upload_forces may follow).
 * `records` may be any host memory; from ls1g_host_alloc memory the copy runs at the full PCIe rate. */
int ls1g_upload_records(ls1g_ctx* ctx, size_t n, const void* records, int format, uint64_t* n_kept, uint64_t* max_id);
/* replaces: DomainDecompBase::writeMoleculesToFile -> FullMolecule::writeBinary over the owned molecules
 * (parallel/DomainDecompBase.cpp:505-540, molecules/FullMolecule.cpp:451-474): ICRVQD records of the owned molecules in the order
 * ls1g_download returns them (cell by cell), *n_out = their number; records == NULL only reports the number.  A context that keeps
 * no orientations (exactly one component with one LJ centre at the origin) writes q = (1 0 0 0), D = 0.
 * scale_v, scale_D: the records carry v * scale_v and D * scale_D (scale_v / scale_D, molecules/MoleculeInterface.h:195-213: one multiplication per component;
 * 1.0 leaves the bits as stored).  For a host loop that applies the velocity scaling of a step inside the NEXT step (ls1g_step's
 * beta_trans / beta_rot) and writes a checkpoint in between: VelocityScalingThermostat::apply (VelocityScalingThermostat.cpp:45-80)
 * has already run in the reference when its writer reads the molecules.  The device state is not changed. */
int ls1g_download_records(ls1g_ctx* ctx, double scale_v, double scale_D, size_t capacity, void* records, size_t* n_out);
/* number of molecules held: owned (non-halo cells) and halo copies */
int ls1g_count(ls1g_ctx* ctx, uint64_t* n_owned, uint64_t* n_halo);
/* replaces: ParticleIterator over ALL_CELLS / ONLY_INNER_AND_BOUNDARY + Molecule getters.
 * Molecules come in cell-sorted order; include_halo selects ALL_CELLS. Returns count in *n_out.
 * field_mask 0 copies nothing and only reports the count (capacity is ignored): during an ls1g_run / ls1g_step sequence molecules
 * that left the block stay with it until the next exchange, so ls1g_count may differ from what a download returns afterwards. */
int ls1g_download(ls1g_ctx* ctx, int field_mask, int include_halo, size_t capacity,
                  ls1g_fields* out, size_t* n_out);
/* Raw SLOTS [first, first + count) of the device's cell-sorted storage, 0 <= first + count <= ls1g_slot_count: owned molecules in
 * the order ls1g_download returns them, interleaved with the slots of (possibly already dropped) halo copies, which carry
 * LS1G_HALO_FLAG in `cell` and must be skipped by the caller.  The output pointers address element 0 of the RANGE.  No compaction
 * pass, and a caller can overlap the copy of one chunk with its own work on the previous one (GpuLinkedCells::downloadMirror
 * refills the host cells that way; ParticleIterator semantics stay with ls1g_download). */
#define LS1G_HALO_FLAG 0x80000000u
int ls1g_slot_count(ls1g_ctx* ctx, uint64_t* n_slots);
int ls1g_download_range(ls1g_ctx* ctx, int field_mask, size_t first, size_t count, ls1g_fields* out);

/* --- per-step operators ------------------------------------------------------------------- */
/* replaces: LinkedCells::update (LinkedCells.cpp:243-356): re-bin into cells (device sort) */
int ls1g_update(ls1g_ctx* ctx);
/* replaces: DomainDecompBase::exchangeMolecules (DomainDecompBase.cpp:51-86,174-225,293-348) on a
 * sequential rank grid, DomainDecomposition::balanceAndExchange (DomainDecomposition.cpp:87-103)
 * otherwise: wrap/send leaving molecules, create/receive halo copies, re-bin. */
int ls1g_exchange(ls1g_ctx* ctx);
/* replaces: LinkedCells::updateMoleculeCaches + traverseCells(VectorizedCellProcessor) +
 * Simulation::updateForces/calcFM (LinkedCells.cpp:564-575,1076-1086; VectorizedCellProcessor.cpp:
 * 796-2821; FullMolecule.cpp:551-631).  Outputs the values endTraversal hands to Domain. */
int ls1g_traverse(ls1g_ctx* ctx, double* upot, double* virial);
/* replaces: LinkedCells::deleteOuterParticles (LinkedCells.cpp:611-628) */
int ls1g_delete_outer(ls1g_ctx* ctx);
/* replaces: Leapfrog::eventNewTimestep -> FullMolecule::upd_preF (Leapfrog.cpp:48-68, FullMolecule.cpp:334-364) */
int ls1g_integrate_pre(ls1g_ctx* ctx, double dt);
/* replaces: Leapfrog::eventForcesCalculated -> upd_postF + Domain::setLocal* (Leapfrog.cpp:70-155) */
int ls1g_integrate_post(ls1g_ctx* ctx, double dt_half, double* summv2, double* sumIw2,
                        uint64_t* n, uint64_t* rot_dof);
/* replaces: VelocityScalingThermostat::apply (VelocityScalingThermostat.cpp:45-80), global betas */
int ls1g_scale(ls1g_ctx* ctx, double beta_trans, double beta_rot);
/* replaces: the per-molecule "explosion heuristics" of Domain::calculateGlobalValues (Domain.cpp:254-295), which the reference
 * runs when a global beta falls below MIN_BETA (Domain.h:20): molecules with U_trans > limit_trans_energy (= KINLIMIT_PER_T * T)
 * get v and F scaled by sqrt(limit / U_trans), molecules with U_rot > 3 * rotDOF * temperature get D and M scaled likewise. */
int ls1g_selective_thermostat(ls1g_ctx* ctx, double limit_trans_energy, double temperature);

/* Device-resident time loop: n_steps x { integrate_pre, update+exchange, traverse, delete_outer,
 * integrate_post, [thermostat] } in the order of Simulation::simulate (Simulation.cpp:979-1177).
 * target_temperature <= 0 : NVE (Domain::thermostatOff); else Domain::calculateGlobalValues betas
 * (Domain.cpp:226-247) are applied each step.  records may be NULL, else [n_steps].
 * Requires forces from a previous ls1g_traverse (as prepare_start does, Simulation.cpp:813-844).
 * NOT reproduced here: the "explosion heuristics" branch of calculateGlobalValues (Domain.cpp:254-311, a beta below MIN_BETA) -
 * the device thermostat of this loop always applies the raw betas.  The host-driven ls1g_step, which takes the Domain's own
 * betas, together with ls1g_selective_thermostat reproduces that branch (GpuLinkedCells does; examples/EOX exercises it). */
int ls1g_run(ls1g_ctx* ctx, int n_steps, double dt, double target_temperature, ls1g_step_record* records);

/* One time step for a HOST-DRIVEN loop (the reference's Simulation::simulate calling the container once per step,
 * Simulation.cpp:979-1177):  VelocityScalingThermostat::apply(beta_trans, beta_rot) of the previous step (Simulation.cpp:1108-1136,
 * fused into this step's pre-kick) -> integrate_pre -> update+exchange -> traverse -> delete_outer -> integrate_post.
 * Equals ls1g_scale + ls1g_run(1 step, NVE); the betas stay the caller's (Domain::calculateGlobalValues).
 * flags: LS1G_STEP_VI = also produce the per-molecule virial Vi and the pair count (needed only when Vi is read back). */
enum { LS1G_STEP_VI = 1 };
int ls1g_step(ls1g_ctx* ctx, double dt, double beta_trans, double beta_rot, int flags, ls1g_step_record* record);

/* Page-locked host staging memory: ls1g_upload / ls1g_download run at full PCIe rate from/to it (any other host pointer
 * works too, at the pageable rate).  replaces: nothing in the reference - it is the device boundary's own buffer. */
int ls1g_host_alloc(ls1g_ctx* ctx, size_t bytes, void** out);
int ls1g_host_free(ls1g_ctx* ctx, void* p);

/* --- multi-GPU (one context per rank/GPU) --------------------------------------------------
 * replaces: DomainDecomposition::getBoundingBoxMin/Max (DomainDecomposition.cpp:116-125); host only */
int ls1g_decomp_box(const int rank_grid[3], const int rank_coords[3], const double global_length[3],
                    double box_min[3], double box_max[3]);
/* replaces: MPI_Dims_create + MPI_Cart_coords (DomainDecomposition.cpp:35-39); host only */
int ls1g_decomp_grid(int n_ranks, int rank, int rank_grid[3], int rank_coords[3]);
/* NCCL bootstrap: rank 0 calls ls1g_comm_unique_id, distributes the 128 bytes, all call ls1g_comm_init.
 * replaces: MPI_Cart_create + CommunicationPartner set-up (DomainDecomposition.cpp:36, DomainDecompMPIBase.cpp:148-170) */
int ls1g_comm_unique_id(unsigned char id[128]);
int ls1g_comm_init(ls1g_ctx* ctx, int n_ranks, int rank, const unsigned char id[128]);
/* replaces: Domain::calculateGlobalValues collComm allreduce (Domain.cpp:170-176,214-224): sums up to 8 doubles in place */
int ls1g_allreduce_sum(ls1g_ctx* ctx, double* values, int n);
/* In-process transport: all ranks are contexts of THIS process (ordered by rank), records move by device copies.
 * Exercises the same pack / stage / unpack kernels as the NCCL path on any number of GPUs >= 1 (tests).
 * replaces: the sequential fall-back of the reference when MPI is absent (DomainDecompBase) only in spirit;
 * the exchanged regions are those of NeighbourCommunicationScheme.cpp:500-600. */
int ls1g_comm_init_local(ls1g_ctx** ctxs, int n);
/* one leaving+halo exchange of all local ranks in lockstep (= ls1g_exchange on every rank) */
int ls1g_exchange_group(ls1g_ctx** ctxs, int n);
/* ls1g_run / ls1g_step for all ranks of a local group in lockstep (one host thread enqueues for every context; the contexts may live
 * on different GPUs of the box).  The same loop as a NCCL rank runs - neighbour lists with persistent halo lists, halo refresh
 * overlapped with the inner blocks of the force pass, guarded speculative steps - with device-to-device copies ordered by events in
 * place of ncclSend/ncclRecv.  records / record hold the SUMS over the ranks, i.e. what one MPI rank owning the whole box would hand
 * to Domain::setLocalUpot etc.
 * replaces: Simulation::simulate on every rank of DomainDecomposition (Simulation.cpp:979-1177, DomainDecomposition.cpp:87-125) */
int ls1g_run_group(ls1g_ctx** ctxs, int n, int n_steps, double dt, double target_temperature, ls1g_step_record* records);
int ls1g_step_group(ls1g_ctx** ctxs, int n, double dt, double beta_trans, double beta_rot, int flags, ls1g_step_record* record);

/* --- measurement -------------------------------------------------------------------------- */
/* LS1G_T_FORCE brackets ONLY the pair kernel (the roofline kernel); its reductions and the site cache are LS1G_T_AUX */
enum { LS1G_T_FORCE = 0, LS1G_T_SORT = 1, LS1G_T_INTEGRATE = 2, LS1G_T_EXCHANGE = 3, LS1G_T_AUX = 4, LS1G_T_COUNT = 5 };
/* accumulated CUDA-event time (ms) and launch count per kernel class since the last reset */
int ls1g_timers(ls1g_ctx* ctx, double ms[LS1G_T_COUNT], uint64_t launches[LS1G_T_COUNT], int reset);
/* the cudaStream_t all work of this context is issued on (for CUDA-event timing by the caller) */
int ls1g_stream(ls1g_ctx* ctx, void** stream);
/* dependent-free DFMA micro-kernel: measured FP64 FMA peak of this GPU in TFLOP/s (roofline denominator) */
int ls1g_measure_fp64_peak(int device, double* tflops);
/* replaces: FlopCounter as the cell processor of one traversal (adapter/FlopCounter.cpp:196-421; count -> FLOP model FlopCounter.h:150-283,
 * multipliers FlopCounter.cpp:19-75), as io/FlopRateWriter.cpp runs it: the reference's own count of the ALGORITHMIC (Newton-3) floating
 * point work of a force pass over the current cells.  Potentials in the order of FlopCounter.h:138-147: LJ, charge, charge-dipole,
 * dipole, charge-quadrupole, dipole-quadrupole, quadrupole.  Needs halo copies in place (after ls1g_exchange, before ls1g_delete_outer). */
typedef struct {
	double molecule_distances;   /* _moleculeDistances */
	double kernel_calls[7];      /* _numKernelCalls per potential */
	double macro_calls[7];       /* _numMacroCalls per potential */
	double flops;                /* _Counts::getTotalFlops() */
} ls1g_flop_counts;
int ls1g_flop_count(ls1g_ctx* ctx, ls1g_flop_counts* out);
/* replaces: RDFCellProcessor as the cell processor of one traversal (adapter/RDFCellProcessor.cpp:18-133) feeding the counters of the
 * RDF plugin (io/RDF.h:95-160: observeRDFMolecule, observeRDFSite, observeARDFMolecule; MoleculeInterface.h:157-190), as
 * RDF::afterForces runs it (io/RDF.cpp:573-581).  The counts of this traversal are ADDED to the caller's arrays (RDF::_distribution.local
 * etc. accumulate over the sampling steps in the same way):
 *   mol  [nc (nc + 1) / 2][bins]           component pairs (i, j >= i) in the order (0,0) (0,1) ... (1,1) ...   = _distribution.local[i][j-i]
 *   site [S][S][bins]                       S = all sites of all components, rows in component order, inside a component LJ centres,
 *                                           charges, dipoles, quadrupoles (Component::numSites order); entry (site0_i + m, site0_j + n) =
 *                                           _siteDistribution.local[i][j-i][m][n]; may be NULL when no component pair has more than 2 sites
 *   ardf [nc][nc][bins * angular_bins]      = _ARDFdistribution.local[i][j]; may be NULL when angular_bins <= 1
 * Needs halo copies in place (after ls1g_exchange, before ls1g_delete_outer).  On a multi-rank context every rank counts the pairs
 * the reference counts on that rank; the caller sums over ranks (RDF::collectRDF). */
typedef struct {
	uint32_t bins;               /* RDF::_bins */
	uint32_t angular_bins;       /* RDF::_angularBins; <= 1: no angular RDF */
	double   interval_length;    /* RDF::_intervalLength; pairs beyond interval_length * bins are not counted */
} ls1g_rdf_params;
int ls1g_rdf(ls1g_ctx* ctx, const ls1g_rdf_params* params, uint64_t* mol, uint64_t* site, uint64_t* ardf);
/* neighbour-list bookkeeping of the device-resident loop (measurement): list builds so far, cell rows per block of the list kernels
 * (4 = 2 x 2 bundles, 1 = single rows), list capacity per molecule, skin.  Any pointer may be NULL. */
int ls1g_list_stats(ls1g_ctx* ctx, uint64_t* rebuilds, int* rows_per_block, int* entries_per_molecule, double* skin);
/* pair statistics of the last traverse: molecule pairs inside the cutoff (counted once) */
int ls1g_last_pair_count(ls1g_ctx* ctx, uint64_t* pairs_in_cutoff);

#ifdef __cplusplus
}
#endif
#endif /* LS1G_H */
