// TEST INFRASTRUCTURE — not product code.
//
// ref_driver: links the UNMODIFIED reference objects (oracle/_ref/libmardyn_ref.a,
// built from /root/reference/src by oracle/ref/Makefile) and dumps what the
// reference computes on a phase-space file, so that the CPU restatement in
// oracle/ls1_oracle.c and the CUDA path can be pinned to it.  The set-up mirrors
// the reference's own test factory
// (src/particleContainer/tests/ParticleContainerFactory.cpp:52-102) and the step
// order of Simulation::simulate (src/Simulation.cpp:979-1177).
//
// Output: a flat little-endian "named array" file (see put()), parsed by
// tests/golden/make_golden.py.
//
//   ref_driver run <file.inp> <rc> <rcLJ> <open|periodic> <ctor|rebuild> <steps> <dt> <out.bin> [epsRF]
//   ref_driver cellindex <L0> <L1> <L2> <rc> <ctor|rebuild> <points.bin> <out.bin>
//   ref_driver rdf <file.inp> <rc> <open|periodic> <ctor|rebuild> <bins> <intervallength> <angularbins> <out.bin>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <string>
#include <vector>
#include <map>
#include <algorithm>

#include "Simulation.h"
#include "Domain.h"
#include "ensemble/EnsembleBase.h"
#include "molecules/Component.h"
#include "molecules/Molecule.h"
#include "io/ASCIIReader.h"
#include "parallel/DomainDecompBase.h"
#include "particleContainer/LinkedCells.h"
#include "particleContainer/adapter/CellProcessor.h"
#include "particleContainer/adapter/VectorizedCellProcessor.h"
#include "particleContainer/adapter/LegacyCellProcessor.h"
#include "particleContainer/adapter/ParticlePairs2PotForceAdapter.h"
#include "particleContainer/adapter/FlopCounter.h"
#include "integrators/Leapfrog.h"
#include "io/RDF.h"
#include "particleContainer/adapter/RDFCellProcessor.h"
#include "utils/Logger.h"

using Log::global_log;

// ---------------------------------------------------------------- output ----
static FILE* g_out = nullptr;

// record: u32 name_len, name, u8 dtype ('d' f64, 'q' i64), u32 ndim, u64 shape[ndim], raw data
static void put(const std::string& name, char dtype, const std::vector<uint64_t>& shape, const void* data) {
	uint32_t nl = name.size();
	fwrite(&nl, 4, 1, g_out);
	fwrite(name.data(), 1, nl, g_out);
	fwrite(&dtype, 1, 1, g_out);
	uint32_t nd = shape.size();
	fwrite(&nd, 4, 1, g_out);
	uint64_t n = 1;
	for (auto s : shape) { fwrite(&s, 8, 1, g_out); n *= s; }
	if (n) fwrite(data, 8, n, g_out);
}
static void putd(const std::string& name, const std::vector<double>& v, std::vector<uint64_t> shape = {}) {
	if (shape.empty()) shape = {v.size()};
	put(name, 'd', shape, v.data());
}
static void puti(const std::string& name, const std::vector<int64_t>& v, std::vector<uint64_t> shape = {}) {
	if (shape.empty()) shape = {v.size()};
	put(name, 'q', shape, v.data());
}

// ------------------------------------------------- recording cell processor --
// Same idea as the reference's CellProcessorStub (LinkedCellsTest.cpp:164-237):
// records every processCell / processCellPair call the C08 traversal emits.
class RecordingCellProcessor : public CellProcessor {
public:
	RecordingCellProcessor() : CellProcessor(0.0, 0.0) {}
	std::vector<int64_t> calls;  // triplets (cell1, cell2, kind) ; kind 0 = processCell, 1 = processCellPair
	void initTraversal() override {}
	void preprocessCell(ParticleCell&) override {}
	void processCellPair(ParticleCell& c1, ParticleCell& c2, bool) override {
		#pragma omp critical(rec)
		{ calls.push_back(c1.getCellIndex()); calls.push_back(c2.getCellIndex()); calls.push_back(1); }
	}
	void processCell(ParticleCell& c) override {
		#pragma omp critical(rec)
		{ calls.push_back(c.getCellIndex()); calls.push_back(c.getCellIndex()); calls.push_back(0); }
	}
	double processSingleMolecule(Molecule*, ParticleCell&) override { return 0.0; }
	void postprocessCell(ParticleCell&) override {}
	void endTraversal() override {}
};

// ------------------------------------------------------------- dump helpers --
static void dumpComponents(Domain* domain) {
	std::vector<Component>& comps = *(global_simulation->getEnsemble()->getComponents());
	std::vector<int64_t> counts;   // per component: nLJ nQ nD nQuad rotdof lookupid
	std::vector<double> scal;      // per component: m I11 I22 I33
	std::vector<double> lj, ch, di, qu;  // lj: x y z m eps sigma shift6 ; ch: x y z m q ; di/qu: x y z ex ey ez abs
	for (auto& c : comps) {
		counts.insert(counts.end(), {(int64_t)c.numLJcenters(), (int64_t)c.numCharges(), (int64_t)c.numDipoles(),
			(int64_t)c.numQuadrupoles(), (int64_t)c.getRotationalDegreesOfFreedom(), (int64_t)c.getLookUpId()});
		scal.insert(scal.end(), {c.m(), c.I11(), c.I22(), c.I33()});
		for (unsigned i = 0; i < c.numLJcenters(); ++i) { auto& s = c.ljcenter(i);
			lj.insert(lj.end(), {s.rx(), s.ry(), s.rz(), s.m(), s.eps(), s.sigma(), s.shift6()}); }
		for (unsigned i = 0; i < c.numCharges(); ++i) { auto& s = c.charge(i);
			ch.insert(ch.end(), {s.rx(), s.ry(), s.rz(), s.m(), s.q()}); }
		for (unsigned i = 0; i < c.numDipoles(); ++i) { auto& s = c.dipole(i);
			di.insert(di.end(), {s.rx(), s.ry(), s.rz(), s.ex(), s.ey(), s.ez(), s.absMy()}); }
		for (unsigned i = 0; i < c.numQuadrupoles(); ++i) { auto& s = c.quadrupole(i);
			qu.insert(qu.end(), {s.rx(), s.ry(), s.rz(), s.ex(), s.ey(), s.ez(), s.absQ()}); }
	}
	puti("comp_counts", counts, {comps.size(), 6});
	putd("comp_scalars", scal, {comps.size(), 4});
	putd("lj_sites", lj, {lj.size() / 7, 7});
	putd("charge_sites", ch, {ch.size() / 5, 5});
	putd("dipole_sites", di, {di.size() / 7, 7});
	putd("quadrupole_sites", qu, {qu.size() / 7, 7});
	putd("mixcoeff", domain->getmixcoeff());
	putd("epsRF", std::vector<double>{domain->getepsilonRF()});
	putd("box", std::vector<double>{domain->getGlobalLength(0), domain->getGlobalLength(1), domain->getGlobalLength(2)});
}

struct MolRow { int64_t id, cid, cell; double r[3], v[3], q[4], L[3], F[3], M[3], Vi[3]; };

// all molecules (incl. halo copies) in container iteration order, single thread
static std::vector<MolRow> snapshot(LinkedCells* lc) {
	std::vector<MolRow> rows;
	for (auto m = lc->iterator(ParticleIterator::ALL_CELLS); m.isValid(); ++m) {
		MolRow w;
		w.id = m->getID(); w.cid = m->componentid();
		w.cell = lc->getCellIndexOfMolecule(&*m);
		for (int d = 0; d < 3; ++d) { w.r[d] = m->r(d); w.v[d] = m->v(d); w.L[d] = m->D(d);
			w.F[d] = m->F(d); w.M[d] = m->M(d); w.Vi[d] = m->Vi(d); }
		w.q[0] = m->q().qw(); w.q[1] = m->q().qx(); w.q[2] = m->q().qy(); w.q[3] = m->q().qz();
		rows.push_back(w);
	}
	return rows;
}

static void dumpRows(const std::string& prefix, const std::vector<MolRow>& rows, bool withForces) {
	size_t n = rows.size();
	std::vector<int64_t> id(n), cid(n), cell(n);
	std::vector<double> r(3 * n), v(3 * n), q(4 * n), L(3 * n), F(3 * n), M(3 * n), Vi(3 * n);
	for (size_t i = 0; i < n; ++i) {
		id[i] = rows[i].id; cid[i] = rows[i].cid; cell[i] = rows[i].cell;
		for (int d = 0; d < 3; ++d) { r[3*i+d] = rows[i].r[d]; v[3*i+d] = rows[i].v[d]; L[3*i+d] = rows[i].L[d];
			F[3*i+d] = rows[i].F[d]; M[3*i+d] = rows[i].M[d]; Vi[3*i+d] = rows[i].Vi[d]; }
		for (int d = 0; d < 4; ++d) q[4*i+d] = rows[i].q[d];
	}
	puti(prefix + "id", id); puti(prefix + "cid", cid); puti(prefix + "cell", cell);
	putd(prefix + "r", r, {n, 3}); putd(prefix + "v", v, {n, 3}); putd(prefix + "q", q, {n, 4}); putd(prefix + "L", L, {n, 3});
	if (withForces) { putd(prefix + "F", F, {n, 3}); putd(prefix + "M", M, {n, 3}); putd(prefix + "Vi", Vi, {n, 3}); }
}

// ------------------------------------------------------------------- run ----
static LinkedCells* makeContainer(double bmin[3], double bmax[3], double rc, const std::string& how) {
	if (how == "ctor") return new LinkedCells(bmin, bmax, rc);       // test-factory path, double rc (LinkedCells.cpp:42-107)
	LinkedCells* lc = new LinkedCells();                               // XML path (Simulation.cpp:422-450), float rc
	lc->setCutoff(rc);
	lc->rebuild(bmin, bmax);
	return lc;
}

static int cmdRun(int argc, char** argv) {
	if (argc < 10) { fprintf(stderr, "usage: run inp rc rcLJ open|periodic ctor|rebuild steps dt out [epsRF]\n"); return 2; }
	std::string inp = argv[2];
	double rc = atof(argv[3]), rcLJ = atof(argv[4]);
	bool periodic = std::string(argv[5]) == "periodic";
	std::string how = argv[6];
	int steps = atoi(argv[7]);
	double dt = atof(argv[8]);
	g_out = fopen(argv[9], "wb");
	double epsRF = argc > 10 ? atof(argv[10]) : 1e10;
	bool legacy = getenv("REF_CP") && std::string(getenv("REF_CP")) == "legacy";

	Simulation* sim = new Simulation();
	Domain* domain = sim->getDomain();
	DomainDecompBase* decomp = &sim->domainDecomposition();
	sim->setcutoffRadius(rc);
	sim->setLJCutoff(rcLJ);
	domain->setepsilonRF(epsRF);

	ASCIIReader reader;
	reader.setPhaseSpaceHeaderFile(inp);
	reader.setPhaseSpaceFile(inp);
	reader.readPhaseSpaceHeader(domain, dt);
	double bmin[3], bmax[3];
	for (int d = 0; d < 3; ++d) { bmin[d] = decomp->getBoundingBoxMin(d, domain); bmax[d] = decomp->getBoundingBoxMax(d, domain); }
	LinkedCells* lc = makeContainer(bmin, bmax, rc, how);
	reader.readPhaseSpace(lc, domain, decomp);
	lc->deleteOuterParticles();
	lc->update();
	lc->updateMoleculeCaches();
	domain->initParameterStreams(rc, rcLJ);

	dumpComponents(domain);
	putd("rc", std::vector<double>{rc, rcLJ});
	putd("dt", std::vector<double>{dt});
	dumpRows("in_", snapshot(lc), false);

	if (periodic) { decomp->exchangeMolecules(lc, domain); lc->updateMoleculeCaches(); }
	{
		int* bw = lc->getBoxWidthInNumCells();  // without halo; one halo layer each side
		puti("cells_per_dim", std::vector<int64_t>{bw[0] + 2, bw[1] + 2, bw[2] + 2});
		putd("cell_length", std::vector<double>{lc->getCellLength()[0], lc->getCellLength()[1], lc->getCellLength()[2]});
	}
	dumpRows("pre_", snapshot(lc), false);

	{   // cell-pair enumeration of the C08 traversal
		RecordingCellProcessor rec;
		lc->traverseCells(rec);
		puti("cellpairs", rec.calls, {rec.calls.size() / 3, 3});
	}
	{   // FLOP model counts (FlopCounter.cpp:19-75)
		FlopCounter fc(rc, rcLJ);
		lc->traverseCells(fc);
		putd("flops", std::vector<double>{fc.getTotalFlopCount()});
	}

	CellProcessor* cp;
	ParticlePairs2PotForceAdapter* adapter = nullptr;
	if (legacy) { adapter = new ParticlePairs2PotForceAdapter(*domain); cp = new LegacyCellProcessor(rc, rcLJ, adapter); }
	else cp = new VectorizedCellProcessor(*domain, rc, rcLJ);

	lc->traverseCells(*cp);
	for (auto m = lc->iterator(ParticleIterator::ALL_CELLS); m.isValid(); ++m) m->calcFM();
	dumpRows("f_", snapshot(lc), true);
	putd("upot_virial", std::vector<double>{domain->getLocalUpot(), domain->getLocalVirial()});

	if (steps > 0) {
		domain->thermostatOff();
		Leapfrog integ(dt);
		integ.init();
		lc->deleteOuterParticles();
		std::vector<double> series;  // per step: upot virial summv2 sumIw2
		for (int s = 0; s < steps; ++s) {
			integ.eventNewTimestep(lc, domain);
			lc->update();
			if (periodic) decomp->exchangeMolecules(lc, domain);
			lc->updateMoleculeCaches();
			lc->traverseCells(*cp);
			for (auto m = lc->iterator(ParticleIterator::ALL_CELLS); m.isValid(); ++m) m->calcFM();
			lc->deleteOuterParticles();
			integ.eventForcesCalculated(lc, domain);
			double summv2 = 0, sumIw2 = 0;
			for (auto m = lc->iterator(ParticleIterator::ONLY_INNER_AND_BOUNDARY); m.isValid(); ++m) {
				summv2 += m->mass() * m->v2();
				sumIw2 += 2.0 * m->U_rot();
			}
			series.insert(series.end(), {domain->getLocalUpot(), domain->getLocalVirial(), summv2, sumIw2});
		}
		putd("series", series, {(uint64_t)steps, 4});
		std::vector<MolRow> rows = snapshot(lc);
		std::sort(rows.begin(), rows.end(), [](const MolRow& a, const MolRow& b) { return a.id < b.id; });
		dumpRows("post_", rows, true);
	}
	fclose(g_out);
	return 0;
}

// ------------------------------------------------------------ cell index ----
static int cmdCellIndex(int argc, char** argv) {
	if (argc < 9) { fprintf(stderr, "usage: cellindex L0 L1 L2 rc ctor|rebuild points.bin out.bin\n"); return 2; }
	double bmin[3] = {0, 0, 0}, bmax[3] = {atof(argv[2]), atof(argv[3]), atof(argv[4])};
	double rc = atof(argv[5]);
	std::string how = argv[6];
	FILE* f = fopen(argv[7], "rb");
	fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
	std::vector<double> pts(sz / 8);
	if (fread(pts.data(), 8, pts.size(), f) != pts.size()) return 3;
	fclose(f);
	g_out = fopen(argv[8], "wb");
	new Simulation();
	LinkedCells* lc = makeContainer(bmin, bmax, rc, how);
	size_t n = pts.size() / 3;
	std::vector<int64_t> idx(n);
	for (size_t i = 0; i < n; ++i) idx[i] = lc->getCellIndexOfPoint(&pts[3 * i]);
	int* bw = lc->getBoxWidthInNumCells();
	puti("cells_per_dim", std::vector<int64_t>{bw[0] + 2, bw[1] + 2, bw[2] + 2});
	putd("cell_length", std::vector<double>{lc->getCellLength()[0], lc->getCellLength()[1], lc->getCellLength()[2]});
	puti("cell_index", idx);
	fclose(g_out);
	return 0;
}

// ------------------------------------------------------------------- rdf ----
// RDF as the reference's own unit test drives it (src/io/tests/RDFTest.cpp:41-90): the plugin object is configured through
// the friend declaration `friend class RDFTest` (src/io/RDF.h:39), one traversal with RDFCellProcessor, local counters dumped.
class RDFTest {
public:
	static int run(int argc, char** argv) {
		if (argc < 10) { fprintf(stderr, "usage: rdf inp rc open|periodic ctor|rebuild bins intervallength angularbins out\n"); return 2; }
		std::string inp = argv[2];
		double rc = atof(argv[3]);
		bool periodic = std::string(argv[4]) == "periodic";
		std::string how = argv[5];
		unsigned long bins = strtoul(argv[6], nullptr, 10);
		double interval = atof(argv[7]);
		unsigned long abins = strtoul(argv[8], nullptr, 10);
		g_out = fopen(argv[9], "wb");

		Simulation* sim = new Simulation();
		Domain* domain = sim->getDomain();
		DomainDecompBase* decomp = &sim->domainDecomposition();
		sim->setcutoffRadius(rc);
		sim->setLJCutoff(rc);
		ASCIIReader reader;
		reader.setPhaseSpaceHeaderFile(inp);
		reader.setPhaseSpaceFile(inp);
		reader.readPhaseSpaceHeader(domain, 0.001);
		double bmin[3], bmax[3];
		for (int d = 0; d < 3; ++d) { bmin[d] = decomp->getBoundingBoxMin(d, domain); bmax[d] = decomp->getBoundingBoxMax(d, domain); }
		LinkedCells* lc = makeContainer(bmin, bmax, rc, how);
		reader.readPhaseSpace(lc, domain, decomp);
		lc->deleteOuterParticles();
		lc->update();
		lc->updateMoleculeCaches();
		dumpComponents(domain);
		putd("rc", std::vector<double>{rc, rc});
		if (periodic) { decomp->exchangeMolecules(lc, domain); lc->updateMoleculeCaches(); }
		{
			int* bw = lc->getBoxWidthInNumCells();
			puti("cells_per_dim", std::vector<int64_t>{bw[0] + 2, bw[1] + 2, bw[2] + 2});
		}
		dumpRows("pre_", snapshot(lc), false);

		std::vector<Component>* comps = global_simulation->getEnsemble()->getComponents();
		RDF rdf;
		rdf._intervalLength = interval;
		rdf._bins = bins;
		rdf._angularBins = abins;
		rdf._components = comps;
		rdf._writeFrequency = 25000;
		rdf._outputPrefix = "out";
		rdf._readConfig = true;
		rdf.init();
		rdf.tickRDF();
		RDFCellProcessor cp(rc, &rdf);
		lc->traverseCells(cp);

		const size_t nC = comps->size();
		std::vector<size_t> site0(nC + 1, 0);
		for (size_t i = 0; i < nC; ++i) site0[i + 1] = site0[i] + (*comps)[i].numSites();
		const size_t S = site0[nC];
		std::vector<int64_t> mol, site(S * S * bins, 0), ardf;
		for (size_t i = 0; i < nC; ++i)
			for (size_t k = 0; i + k < nC; ++k) {
				for (size_t l = 0; l < bins; ++l) mol.push_back((int64_t)rdf._distribution.local[i][k][l]);
				const size_t ni = (*comps)[i].numSites(), nj = (*comps)[i + k].numSites();
				if (ni + nj > 2)
					for (size_t m = 0; m < ni; ++m)
						for (size_t n = 0; n < nj; ++n)
							for (size_t l = 0; l < bins; ++l)
								site[((site0[i] + m) * S + site0[i + k] + n) * bins + l] = (int64_t)rdf._siteDistribution.local[i][k][m][n][l];
			}
		if (rdf._doARDF)
			for (size_t i = 0; i < nC; ++i)
				for (size_t k = 0; k < nC; ++k)
					for (size_t l = 0; l < bins * abins; ++l) ardf.push_back((int64_t)rdf._ARDFdistribution.local[i][k][l]);
		puti("rdf_mol", mol, {mol.size() / bins, bins});
		puti("rdf_site", site, {S, S, bins});
		puti("rdf_ardf", ardf, {ardf.size()});
		putd("rdf_params", std::vector<double>{(double)bins, interval, (double)abins, rdf._doCollectSiteRDF ? 1.0 : 0.0});
		fclose(g_out);
		return 0;
	}
};

int main(int argc, char** argv) {
	global_log = new Log::Logger(Log::Error);
	if (argc < 2) { fprintf(stderr, "usage: ref_driver run|cellindex|rdf ...\n"); return 2; }
	std::string cmd = argv[1];
	if (cmd == "run") return cmdRun(argc, argv);
	if (cmd == "cellindex") return cmdCellIndex(argc, argv);
	if (cmd == "rdf") return RDFTest::run(argc, argv);
	fprintf(stderr, "unknown command %s\n", cmd.c_str());
	return 2;
}
