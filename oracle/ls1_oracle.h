/*
 * TEST INFRASTRUCTURE — NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load this library.
 *
 * ls1_oracle: plain-C, scalar, single-threaded restatement of the reference's short-range force +
 * integrate path (LinkedCells + C08 traversal + VectorizedCellProcessor + calcFM + Leapfrog +
 * DomainDecompBase periodic images).  Each function cites the reference file:line it follows.
 * PINNED: tests/test_oracle_golden.py checks it against tests/golden/ (npz files), which were produced by
 * the unmodified reference itself (oracle/_ref/ref_driver, script tests/golden/make_golden.py).
 */
#ifndef LS1_ORACLE_H
#define LS1_ORACLE_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
	double box_min[3], box_max[3];
	double halo_min[3], halo_max[3];
	double cell_len[3], cell_len_recip[3];
	int    box_cells[3];     /* _boxWidthInNumCells */
	int    cpd[3];           /* _cellsPerDimension (incl. one halo layer each side) */
	double cutoff;
} o_grid;

/* flattened component tables, same layout as the ref_driver dump */
typedef struct {
	int            nc;
	const int64_t* counts;   /* [nc][6] nLJ nCharge nDipole nQuadrupole rotDOF lookUpId */
	const double*  scal;     /* [nc][4] m I11 I22 I33 */
	const double*  lj;       /* [sum nLJ][7] x y z m eps sigma shift6 */
	const double*  ch;       /* [sum nC][5]  x y z m q */
	const double*  di;       /* [sum nD][7]  x y z ex ey ez abs */
	const double*  qu;       /* [sum nQ][7]  x y z ex ey ez abs */
	const double*  mix;      /* (xi, eta) per pair i<j */
	int            nmix;
	double         eps_rf, rc, rc_lj;
} o_params;

/* mode 0 = LinkedCells::rebuild (float rc), 1 = constructor (double rc) */
int      o_grid_init(o_grid* g, const double box_min[3], const double box_max[3], double rc, int mode);
uint32_t o_cell_index(const o_grid* g, const double p[3]);
void     o_cell_indices(const o_grid* g, size_t n, const double* pts, uint32_t* out);
int      o_is_halo_cell(const o_grid* g, uint32_t cell);
/* (cell1, cell2, kind) triplets in the order C08 emits them with one thread; returns count */
size_t   o_c08_pairs(const o_grid* g, uint32_t* triplets, size_t capacity);

/* LJ tables indexed by global LJ-site id: eps24[S*S], sig2[S*S], shift6[S*S]; returns S */
int      o_lj_tables(const o_params* p, double* eps24, double* sig2, double* shift6, int cap);

/* periodic images, DomainDecompBase::exchangeMolecules on one process.
 * In: n molecules (r modified in place by the leaving-particle wrap).  Out: halo copies appended:
 * src index + position.  Returns number of halo copies (or required count if capacity too small). */
size_t   o_exchange(const o_grid* g, const double global_len[3], size_t n, double* r /*[n][3]*/,
                    size_t capacity, uint32_t* halo_src, double* halo_r /*[cap][3]*/);

/* force pass over n molecules (owned + halo copies): normalises q in place (setupSoACache), returns
 * F, M, Vi per molecule (calcFM) and Domain's upot / virial.  stats (may be NULL):
 * [0] molecule-centre distance checks, [1] molecule pairs inside rc (each once). */
int      o_forces(const o_params* p, const o_grid* g, size_t n, const uint32_t* cid, const double* r,
                  double* q, double* F, double* M, double* Vi, double* upot, double* virial,
                  uint64_t* stats);

/* FullMolecule::upd_preF / upd_postF on n molecules of which only those with owned[i]!=0 move */
void     o_upd_preF(const o_params* p, size_t n, const uint32_t* cid, double dt, double* r, double* v,
                    double* q, double* L, const double* F, const double* M);
void     o_upd_postF(const o_params* p, size_t n, const uint32_t* cid, double dt_half, double* v,
                     const double* q, double* L, const double* F, const double* M,
                     double* summv2, double* sumIw2, uint64_t* rot_dof);

/* whole NVE/NVT loop on one periodic process (Simulation::simulate order); state arrays in/out,
 * series[steps][4] = upot, virial, summv2, sumIw2.  target_T<=0: NVE. Returns 0 on success. */
int      o_run(const o_params* p, const o_grid* g, const double global_len[3], size_t n,
               const uint32_t* cid, double* r, double* v, double* q, double* L, double* F, double* M,
               int steps, double dt, double target_T, double* series);

#ifdef __cplusplus
}
#endif
#endif
