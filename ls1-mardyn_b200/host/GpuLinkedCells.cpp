/*
 * GpuLinkedCells.cpp — see GpuLinkedCells.h.  Every device operation is one call of the C ABI (include/ls1g.h); this file
 * only converts between the reference's host objects (Molecule, Component, Domain) and plain arrays.
 */
#include "GpuLinkedCells.h"

#include "Domain.h"
#include "Simulation.h"
#include "ensemble/EnsembleBase.h"
#include "io/RDF.h"
#include "molecules/Component.h"
#include "molecules/Molecule.h"
#include "particleContainer/adapter/CellProcessor.h"
#include "particleContainer/adapter/LegacyCellProcessor.h"
#include "particleContainer/adapter/RDFCellProcessor.h"
#include "particleContainer/adapter/VectorizedCellProcessor.h"
#include "utils/Logger.h"
#include "utils/xmlfileUnits.h"

extern "C" {
#include "ls1g.h"
}

#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <memory>
#include <numeric>
#include <set>
#include <string>
#include <thread>

#if defined(_OPENMP)
#include <omp.h>
#endif

using Log::global_log;

namespace {
double wall() {
	return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}
}  // namespace

GpuLinkedCells::GpuLinkedCells() : LinkedCells() {}

GpuLinkedCells::~GpuLinkedCells() {
	// no logging here: main deletes global_log before the Simulation object (and with it this container) is destroyed
	if (_ctx != nullptr) {
		if (_stage != nullptr) {
			if (_stagePinned) ls1g_host_free(_ctx, _stage);
			else std::free(_stage);
		}
		ls1g_destroy(_ctx);
	}
}

void GpuLinkedCells::printSummary() const {
	double ms[LS1G_T_COUNT];
	uint64_t launches[LS1G_T_COUNT];
	if (_ctx != nullptr && ls1g_timers(_ctx, ms, launches, 0) == 0) {
		global_log->info() << "GpuLinkedCells: kernel time [ms] force " << ms[LS1G_T_FORCE] << " (" << launches[LS1G_T_FORCE] << " launches), sort "
		                   << ms[LS1G_T_SORT] << " (" << launches[LS1G_T_SORT] << "), integrate " << ms[LS1G_T_INTEGRATE] << " ("
		                   << launches[LS1G_T_INTEGRATE] << "), exchange " << ms[LS1G_T_EXCHANGE] << " (" << launches[LS1G_T_EXCHANGE]
		                   << "), aux " << ms[LS1G_T_AUX] << " (" << launches[LS1G_T_AUX] << ")" << std::endl;
	}
	global_log->info() << "GpuLinkedCells: " << _steps << " device steps took " << _tDevice << " s; upload "
	                   << _tUpload << " s (" << _bytesUp << " B); " << _mirrors << " host mirror download(s) took "
	                   << _tMirror << " s (" << _bytesDown << " B); context creation " << _tContext << " s; selective thermostat applied "
	                   << _selective << " time(s); " << _rdfSamples << " RDF sample(s) on the device; first steps [s] " << _tFirstSteps[0] << " " << _tFirstSteps[1] << " "
	                   << _tFirstSteps[2] << " " << _tFirstSteps[3] << ", slowest step " << _tMaxStep
	                   << "; upload = gather " << _tGather << " + copy " << _tH2D << ", mirror = copy " << _tD2H << " + fill " << _tFill << std::endl;
}

void GpuLinkedCells::fail(const char* what) const {
	global_log->error_always_output() << "GpuLinkedCells: " << what << ": " << ls1g_last_error() << std::endl;
	Simulation::exit(1);      // the reference's own fatal path (LinkedCells.cpp:101,162,239,549)
	std::abort();
}

void GpuLinkedCells::chk(int rc, const char* what) const {
	if (rc != 0) fail(what);
}

void GpuLinkedCells::readXML(XMLfileUnits& xmlconfig) {
	LinkedCells::readXML(xmlconfig);
	if (xmlconfig.getNodeValue_int("cellsInCutoffRadius", 1) != 1) {
		global_log->error_always_output() << "GpuLinkedCells: cellsInCutoffRadius must be 1" << std::endl;
		Simulation::exit(1);
	}
	std::string mode = "fused";
	xmlconfig.getNodeValue("mode", mode);
	if (mode == "fused") {
		_mode = FUSED;
	} else if (mode == "operators") {
		_mode = OPERATORS;
	} else {
		global_log->error_always_output() << "GpuLinkedCells: unknown mode " << mode << " (fused | operators)" << std::endl;
		Simulation::exit(1);
	}
	// per-molecule host passes of the time loop that would silently find empty cells in the device phase
	{
		const std::string oldpath = xmlconfig.getcurrentnodepath();
		std::string lrc;
		if (xmlconfig.changecurrentnode("/mardyn/simulation/algorithm/longrange")) xmlconfig.getNodeValue("@type", lrc);
		xmlconfig.changecurrentnode(oldpath);
		const bool tempControl = xmlconfig.query("/mardyn/simulation/algorithm/thermostats/thermostat[@type='TemperatureControl']").card() > 0;
		if (lrc == "planar" || tempControl) {
			global_log->error_always_output() << "GpuLinkedCells: " << (tempControl ? "the TemperatureControl thermostat" : "the planar long-range correction")
			                                  << " works on host molecules inside the time loop and is not supported" << std::endl;
			Simulation::exit(1);
		}
	}
	std::string mirror = "auto";
	xmlconfig.getNodeValue("mirrorfrequency", mirror);
	if (mirror == "auto") {
		_mirrorFrequency = mirrorFrequencyFromPlugins(xmlconfig);
	} else {
		_mirrorFrequency = static_cast<unsigned long>(xmlconfig.getNodeValue_long("mirrorfrequency", 0));
	}
	_device = xmlconfig.getNodeValue_int("device", 0);
	global_log->info() << "GpuLinkedCells: mode " << mode << ", mirror frequency " << _mirrorFrequency
	                   << (mirror == "auto" ? " (from the plugins' write frequencies)" : "") << ", device " << _device << std::endl;
}

// <mirrorfrequency>auto</mirrorfrequency> (the default): the end-of-step plugins that read molecules decide on which steps
// the host cells must hold a snapshot (PluginBase::endStep, Simulation.cpp:1150-1158).  Plugins that only use Domain
// values or timers are ignored; known molecule readers contribute their <writefrequency>; any other plugin is assumed to
// read molecules on every step.  Mirrors are read-only snapshots: a plugin that CHANGES molecules is not supported, and
// the ones that act before or after the force pass (PluginBase::beforeForces/afterForces) never see molecules.
unsigned long GpuLinkedCells::mirrorFrequencyFromPlugins(XMLfileUnits& xmlconfig) const {
	static const std::set<std::string> noMolecules = {"ResultWriter", "SysMonOutput", "TimerWriter", "LoadbalanceWriter",
	                                                  "DecompWriter", "CommunicationPartnerWriter", "KDTreePrinter",
	                                                  "RDF"};      // RDF: its traversal is counted on the device (sampleRDF)
	static const std::set<std::string> readers = {"CheckpointWriter", "MPICheckpointWriter", "XyzWriter", "MmpldWriter",
	                                              "MmspdWriter", "MmspdBinWriter", "PovWriter", "VISWriter", "MaxWriter",
	                                              "EnergyLogWriter"};
	const std::string oldpath = xmlconfig.getcurrentnodepath();
	unsigned long freq = 0;
	for (const char* category : {"/mardyn/simulation/output/outputplugin", "/mardyn/simulation/plugin"}) {
		XMLfile::Query query = xmlconfig.query(category);
		for (auto it = query.begin(); it; ++it) {
			xmlconfig.changecurrentnode(it);
			std::string name;
			xmlconfig.getNodeValue("@name", name);
			bool enabled = true;
			xmlconfig.getNodeValue("@enabled", enabled);
			if (!enabled || noMolecules.count(name) != 0) continue;
			unsigned long f = 1;
			if (readers.count(name) != 0) {
				const long w = xmlconfig.getNodeValue_long(name == "MmpldWriter" ? "writecontrol/writefrequency" : "writefrequency", 1);
				f = w > 0 ? static_cast<unsigned long>(w) : 1;
			} else {
				global_log->warning() << "GpuLinkedCells: plugin " << name << " is not known to be a read-only end-of-step writer: "
				                      << "the host cells are refilled on every step for it, changes it makes to molecules are NOT "
				                      << "carried to the device" << std::endl;
			}
			global_log->info() << "GpuLinkedCells: plugin " << name << " reads molecules every " << f << " step(s)" << std::endl;
			freq = std::gcd(freq, f);
		}
	}
	xmlconfig.changecurrentnode(oldpath);
	return freq;
}

bool GpuLinkedCells::rebuild(double bBoxMin[3], double bBoxMax[3]) {
	if (_ctx != nullptr) {      // new bounding box: the device context is rebuilt at its next use
		if (_devicePhase && !_mirrorValid) downloadMirror();
		if (_stage != nullptr) {
			if (_stagePinned) ls1g_host_free(_ctx, _stage);
			else std::free(_stage);
		}
		_stage = nullptr;
		_stageMolecules = 0;
		ls1g_destroy(_ctx);
		_ctx = nullptr;
		_deviceLoaded = false;
		_devicePhase = false;
		_mirrorValid = false;
	}
	return LinkedCells::rebuild(bBoxMin, bBoxMax);
}

// The context is created at first use, not in rebuild(): components, mixing coefficients and the reaction-field epsilon
// are only complete once Simulation::readXML has finished.
void GpuLinkedCells::ensureContext() {
	if (_ctx != nullptr) return;
	const double t0 = wall();
	Domain* domain = global_simulation->getDomain();
	// component-wise and directed-velocity thermostats act per molecule on the host (VelocityScalingThermostat.cpp:45-80)
	if (domain->severalThermostats() || domain->thermostatIsUndirected(0)) {
		global_log->error_always_output() << "GpuLinkedCells: component-wise and undirected thermostats are not supported" << std::endl;
		Simulation::exit(1);
	}
	ls1g_config cfg{};
	cfg.device = _device;
	for (int d = 0; d < 3; ++d) {
		cfg.box_min[d] = getBoundingBoxMin(d);
		cfg.box_max[d] = getBoundingBoxMax(d);
		cfg.global_length[d] = domain->getGlobalLength(d);
		cfg.rank_grid[d] = 1;
		cfg.rank_coords[d] = 0;
	}
	cfg.cutoff = getCutoff();
	cfg.lj_cutoff = global_simulation->getLJCutoff();
	cfg.eps_rf = domain->getepsilonRF();
	cfg.grid_mode = LS1G_GRID_REBUILD;
	cfg.capacity = 0;
	chk(ls1g_create(&cfg, &_ctx), "ls1g_create");
	uploadComponents();
	int cpd[3];
	double len[3];
	chk(ls1g_get_grid(_ctx, cpd, len), "ls1g_get_grid");
	const double* mine = getCellLength();
	for (int d = 0; d < 3; ++d) {
		if (len[d] != mine[d]) {
			global_log->error_always_output() << "GpuLinkedCells: device cell grid differs from LinkedCells::rebuild in dimension "
			                                  << d << " (" << len[d] << " vs " << mine[d] << ")" << std::endl;
			Simulation::exit(1);
		}
	}
	_numCells = static_cast<size_t>(cpd[0]) * cpd[1] * cpd[2];
	global_log->info() << "GpuLinkedCells: device context created, " << cpd[0] << " x " << cpd[1] << " x " << cpd[2]
	                   << " cells" << std::endl;
	_tContext += wall() - t0;
}

// Component tables (src/molecules/Component.h, Site.h) and Domain::getmixcoeff -> ls1g_set_components
void GpuLinkedCells::uploadComponents() {
	std::vector<Component>& comps = *global_simulation->getEnsemble()->getComponents();
	const size_t nc = comps.size();
	std::vector<std::vector<double>> lj(nc), ch(nc), di(nc), qu(nc);
	std::vector<ls1g_component> rows(nc);
	for (size_t k = 0; k < nc; ++k) {
		const Component& c = comps[k];
		for (unsigned i = 0; i < c.numLJcenters(); ++i) {
			const LJcenter& s = c.ljcenter(i);
			lj[k].insert(lj[k].end(), {s.rx(), s.ry(), s.rz(), s.m(), s.eps(), s.sigma(), s.shift6()});
		}
		for (unsigned i = 0; i < c.numCharges(); ++i) {
			const Charge& s = c.charge(i);
			ch[k].insert(ch[k].end(), {s.rx(), s.ry(), s.rz(), s.m(), s.q()});
		}
		for (unsigned i = 0; i < c.numDipoles(); ++i) {
			const Dipole& s = c.dipole(i);
			di[k].insert(di[k].end(), {s.rx(), s.ry(), s.rz(), s.ex(), s.ey(), s.ez(), s.absMy()});
		}
		for (unsigned i = 0; i < c.numQuadrupoles(); ++i) {
			const Quadrupole& s = c.quadrupole(i);
			qu[k].insert(qu[k].end(), {s.rx(), s.ry(), s.rz(), s.ex(), s.ey(), s.ez(), s.absQ()});
		}
		ls1g_component& r = rows[k];
		r.mass = c.m();
		r.inertia[0] = c.I11();
		r.inertia[1] = c.I22();
		r.inertia[2] = c.I33();
		r.rot_dof = static_cast<int>(c.getRotationalDegreesOfFreedom());
		r.n_lj = static_cast<int>(c.numLJcenters());
		r.n_charge = static_cast<int>(c.numCharges());
		r.n_dipole = static_cast<int>(c.numDipoles());
		r.n_quadrupole = static_cast<int>(c.numQuadrupoles());
		r.lj = lj[k].empty() ? nullptr : lj[k].data();
		r.charge = ch[k].empty() ? nullptr : ch[k].data();
		r.dipole = di[k].empty() ? nullptr : di[k].data();
		r.quadrupole = qu[k].empty() ? nullptr : qu[k].data();
	}
	// the library's own rule for "no orientation kept" (one component, one LJ centre at the origin, no rotational DOF)
	_pointMolecules = nc == 1 && rows[0].n_lj == 1 && rows[0].n_charge == 0 && rows[0].n_dipole == 0 && rows[0].n_quadrupole == 0 &&
	                  rows[0].rot_dof == 0 && lj[0][0] == 0. && lj[0][1] == 0. && lj[0][2] == 0.;
	std::vector<double>& mix = global_simulation->getDomain()->getmixcoeff();
	chk(ls1g_set_components(_ctx, static_cast<int>(nc), rows.data(), mix.empty() ? nullptr : mix.data(),
	                        static_cast<int>(mix.size())),
	    "ls1g_set_components");
}

// One page-locked block (ls1g_host_alloc) carved into the arrays of ls1g_upload / ls1g_download; allocated once, grown on demand.
double GpuLinkedCells::ensureStaging(size_t molecules) {
	if (_stage != nullptr && _stageMolecules >= molecules) return 0.;
	const double t0 = wall();
	const char* pinnedEnv = std::getenv("LS1G_SHIM_PINNED");       // 0: plain (pageable) host memory for the staging block
	const bool pinned = !(pinnedEnv != nullptr && std::atoi(pinnedEnv) == 0);
	if (_stage != nullptr) {
		if (_stagePinned) chk(ls1g_host_free(_ctx, _stage), "ls1g_host_free");
		else std::free(_stage);
	}
	_stage = nullptr;
	_stageMolecules = molecules + molecules / 16 + 1024;
	_stagePinned = pinned;
	// single-centre molecules keep no q / D / M on the device: their 80 bytes per molecule are not staged (nor page-locked)
	const size_t perMolecule = kStageBytesPerMolecule - (_pointMolecules ? 32 + 24 + 24 : 0);
	if (pinned) {
		chk(ls1g_host_alloc(_ctx, _stageMolecules * perMolecule, &_stage), "ls1g_host_alloc");
	} else {
		_stage = std::malloc(_stageMolecules * perMolecule);
		if (_stage == nullptr) fail("staging allocation");
	}
	char* p = static_cast<char*>(_stage);
	auto take = [&](size_t bytesPerMolecule) { void* q = p; p += _stageMolecules * bytesPerMolecule; return q; };
	_sR = static_cast<double*>(take(24));      // 8-byte aligned arrays first
	_sV = static_cast<double*>(take(24));
	_sQ = _pointMolecules ? nullptr : static_cast<double*>(take(32));
	_sL = _pointMolecules ? nullptr : static_cast<double*>(take(24));
	_sF = static_cast<double*>(take(24));
	_sM = _pointMolecules ? nullptr : static_cast<double*>(take(24));
	_sVi = static_cast<double*>(take(24));
	_sId = static_cast<uint64_t*>(take(8));
	_sCid = static_cast<uint32_t*>(take(4));
	_sCell = static_cast<uint32_t*>(take(4));
	_tContext += wall() - t0;
	return wall() - t0;
}

// host cells (owned molecules) -> ls1g_upload.  The device orders molecules by (cell, id), so the gather order is free;
// here it is cell order: count per cell, prefix sum, then every thread fills the slices of its cells.
void GpuLinkedCells::uploadFromHost() {
	ensureContext();
	const double t0 = wall();
	const long nc = static_cast<long>(_numCells);
	std::vector<size_t> first(nc + 1, 0);
#if defined(_OPENMP)
#pragma omp parallel for schedule(static)
#endif
	for (long c = 0; c < nc; ++c) {
		ParticleCell& pc = getCellReference(static_cast<int>(c));
		first[c + 1] = pc.isHaloCell() ? 0 : static_cast<size_t>(pc.getMoleculeCount());
	}
	for (long c = 0; c < nc; ++c) first[c + 1] += first[c];
	const size_t n = first[nc];
	// set-up (page-locking), accounted with the context creation.  Sized for the mirror, which brings the device's slots back:
	// the owned molecules plus one slot per periodic image in the halo cells (on average owned * halo cells / real cells).
	double tot = 1., real = 1.;
	const int* realCells = getBoxWidthInNumCells();          // without the halo layer (one cell wide: cellsInCutoffRadius == 1)
	for (int d = 0; d < 3; ++d) {
		tot *= realCells[d] + 2;
		real *= std::max(1, realCells[d]);
	}
	const size_t slots = static_cast<size_t>(std::min(28.0 * n + 64., n * (1. + (tot / real - 1.) * 1.3) + 4096.));
	double tStage = ensureStaging(std::max(n, slots));
	if (!_deviceLoaded) {                         // ... and so is the allocation of the device arrays
		const double tr = wall();
		chk(ls1g_reserve(_ctx, n), "ls1g_reserve");
		_tContext += wall() - tr;
		tStage += wall() - tr;
	}
	uint64_t* id = _sId;
	uint32_t* cid = _sCid;
	double *r = _sR, *v = _sV, *q = _sQ, *L = _sL;
#if defined(_OPENMP)
#pragma omp parallel for schedule(static)
#endif
	for (long c = 0; c < nc; ++c) {
		if (first[c + 1] == first[c]) continue;
		size_t i = first[c];
		ParticleCell& pc = getCellReference(static_cast<int>(c));
		for (auto m = pc.iterator(); m.isValid(); ++m, ++i) {
			id[i] = m->getID();
			cid[i] = m->componentid();
			for (int d = 0; d < 3; ++d) {
				r[3 * i + d] = m->r(d);
				v[3 * i + d] = m->v(d);
			}
			if (_pointMolecules) continue;      // the device keeps no orientation for a single LJ centre at the origin
			for (int d = 0; d < 3; ++d) L[3 * i + d] = m->D(d);
			const Quaternion& qq = m->q();
			q[4 * i] = qq.qw();
			q[4 * i + 1] = qq.qx();
			q[4 * i + 2] = qq.qy();
			q[4 * i + 3] = qq.qz();
		}
	}
	const double tCopy0 = wall();
	_tGather += tCopy0 - t0 - tStage;
	chk(ls1g_upload(_ctx, n, id, cid, r, v, _pointMolecules ? nullptr : q, _pointMolecules ? nullptr : L), "ls1g_upload");
	_tH2D += wall() - tCopy0;
	_bytesUp += n * (8 + 4 + (_pointMolecules ? 6 : 13) * 8);
	_deviceLoaded = true;
	_tUpload += wall() - t0 - tStage;
}

// LinkedCells::clear (LinkedCells.cpp:590-595) walks the cells serially; with 10^7 host molecules to free that is the slowest
// host pass of the first time step, so the cells are emptied by all threads.
void GpuLinkedCells::clearHostCells() {
	const long nc = static_cast<long>(_numCells);
	if (nc == 0) {
		LinkedCells::clear();
		return;
	}
#if defined(_OPENMP)
#pragma omp parallel for schedule(static)
#endif
	for (long c = 0; c < nc; ++c) getCellReference(static_cast<int>(c)).deallocateAllParticles();
}

// In the device phase the host cells are empty: hand out iterators that are invalid from the start instead of ones that walk
// every (empty) cell, so that the host-side passes of the time loop cost nothing.
ParticleIterator GpuLinkedCells::iterator(ParticleIterator::Type t) {
	if (_devicePhase && !_mirrorValid) return ParticleIterator();
	return LinkedCells::iterator(t);
}

RegionParticleIterator GpuLinkedCells::regionIterator(const double startRegion[3], const double endRegion[3], ParticleIterator::Type type) {
	if (_devicePhase && !_mirrorValid) return RegionParticleIterator();
	return LinkedCells::regionIterator(startRegion, endRegion, type);
}

void GpuLinkedCells::update() {
	if (!_devicePhase) {
		LinkedCells::update();
		return;
	}
	if (_mode == OPERATORS) {
		const double t0 = wall();
		// wrap leaving molecules, create the periodic halo images, re-sort: LinkedCells::update +
		// DomainDecompBase::exchangeMolecules in one call (the host-side exchange that follows finds empty cells)
		chk(ls1g_exchange(_ctx), "ls1g_exchange");
		_tDevice += wall() - t0;
	}
}

void GpuLinkedCells::updateMoleculeCaches() {
	// host phase: the reference's own SoA caches, so that host passes over the molecules (Simulation::updateForces ->
	// calcFM, plugins) stay valid; device phase: site positions are computed inside the force kernels
	if (!_devicePhase) LinkedCells::updateMoleculeCaches();
}

void GpuLinkedCells::traverseCells(CellProcessor& cellProcessor) {
	const bool forceProcessor = dynamic_cast<VectorizedCellProcessor*>(&cellProcessor) != nullptr ||
	                            dynamic_cast<LegacyCellProcessor*>(&cellProcessor) != nullptr;
	if (!forceProcessor) {
		if (_devicePhase && !_mirrorValid && dynamic_cast<RDFCellProcessor*>(&cellProcessor) != nullptr) {
			sampleRDF();      // RDF::afterForces (io/RDF.cpp:573-581): the pair histograms are counted on the device
			return;
		}
		// other counting / sampling processors (FlopCounter, ...) run on the host cells
		if (_devicePhase && !_mirrorValid) {
			global_log->error_always_output() << "GpuLinkedCells: only the force cell processors run on the device; other cell "
			                                     "processors need the host cells (host phase or a mirror step)" << std::endl;
			Simulation::exit(1);
		}
		LinkedCells::traverseCells(cellProcessor);
		return;
	}
	Domain* domain = global_simulation->getDomain();
	double upot = 0., virial = 0.;
	if (!_devicePhase) {
		// prepare_start (Simulation.cpp:813-844): the host cells are authoritative; the device gets a copy and keeps F
		uploadFromHost();
		const double t0 = wall();
		chk(ls1g_exchange(_ctx), "ls1g_exchange");
		chk(ls1g_traverse(_ctx, &upot, &virial), "ls1g_traverse");
		_tDevice += wall() - t0;
	} else if (_mode == FUSED && _inStep && !_ranThisStep) {
		const double t0 = wall();
		ls1g_step_record rec{};
		// [thermostat of the previous step] -> pre -> exchange -> forces (+calcFM) -> deleteOuter -> post, as one device-resident
		// step; the betas are the Domain's own (integratePre).  Vi and the pair count only when the host will read them.
		chk(ls1g_step(_ctx, _dt, _pendingBt, _pendingBr, mirrorWanted() ? LS1G_STEP_VI : 0, &rec), "ls1g_step");
		_pendingBt = _pendingBr = 1.0;
		upot = rec.upot;
		virial = rec.virial;
		_summv2 = rec.summv2;
		_sumIw2 = rec.sumIw2;
		_n = rec.n;
		_rotDOF = rec.rot_dof;
		_ranThisStep = true;
		const double dtStep = wall() - t0;
		_tDevice += dtStep;
		if (_steps < 4) _tFirstSteps[_steps] = dtStep;
		if (dtStep > _tMaxStep) _tMaxStep = dtStep;
	} else {
		const double t0 = wall();
		chk(ls1g_traverse(_ctx, &upot, &virial), "ls1g_traverse");
		_tDevice += wall() - t0;
	}
	domain->setLocalUpot(upot);          // VectorizedCellProcessor::endTraversal (VectorizedCellProcessor.cpp:155-156)
	domain->setLocalVirial(virial);
}

// io/RDF.h:39 declares `friend class RDFTest` for the reference's unit test (io/tests/RDFTest.cpp, which is not part of the
// executable).  The same door lets the container read the plugin's binning and add the device-side counts of one traversal
// to its counters; upstream, a public RDF::addLocalCounts(...) would take its place.
class RDFTest {
public:
	struct Shape { unsigned long bins, angularBins; double interval; bool site, angular; };
	static Shape shape(const RDF& rdf) {
		return Shape{rdf._bins, rdf._angularBins, rdf._intervalLength, rdf._doCollectSiteRDF, rdf._doARDF};
	}
	static void add(RDF& rdf, const std::vector<Component>& comps, const std::vector<uint64_t>& mol,
	                const std::vector<uint64_t>& site, const std::vector<uint64_t>& ardf) {
		const size_t nC = comps.size(), bins = rdf._bins;
		std::vector<size_t> site0(nC + 1, 0);
		for (size_t i = 0; i < nC; ++i) site0[i + 1] = site0[i] + comps[i].numSites();
		const size_t S = site0[nC];
		size_t row = 0;
		for (size_t i = 0; i < nC; ++i) {
			for (size_t k = 0; i + k < nC; ++k, ++row) {
				for (size_t l = 0; l < bins; ++l) rdf._distribution.local[i][k][l] += mol[row * bins + l];
				const size_t ni = comps[i].numSites(), nj = comps[i + k].numSites();
				if (ni + nj > 2) {
					for (size_t m = 0; m < ni; ++m)
						for (size_t n = 0; n < nj; ++n)
							for (size_t l = 0; l < bins; ++l)
								rdf._siteDistribution.local[i][k][m][n][l] += site[((site0[i] + m) * S + site0[i + k] + n) * bins + l];
				}
			}
			if (rdf._doARDF) {
				const size_t per = bins * rdf._angularBins;
				for (size_t k = 0; k < nC; ++k)
					for (size_t l = 0; l < per; ++l) rdf._ARDFdistribution.local[i][k][l] += ardf[(i * nC + k) * per + l];
			}
		}
	}
};

void GpuLinkedCells::sampleRDF() {
	RDF* rdf = nullptr;
	for (PluginBase* p : *global_simulation->getPluginList()) {
		rdf = dynamic_cast<RDF*>(p);
		if (rdf != nullptr) break;
	}
	if (rdf == nullptr) {
		global_log->error_always_output() << "GpuLinkedCells: RDFCellProcessor without an RDF plugin" << std::endl;
		Simulation::exit(1);
	}
	const double t0 = wall();
	const std::vector<Component>& comps = *global_simulation->getEnsemble()->getComponents();
	const RDFTest::Shape sh = RDFTest::shape(*rdf);
	size_t S = 0;
	for (const Component& c : comps) S += c.numSites();
	const size_t nC = comps.size();
	std::vector<uint64_t> mol(nC * (nC + 1) / 2 * sh.bins, 0), site(sh.site ? S * S * sh.bins : 0, 0),
	    ardf(sh.angular ? nC * nC * sh.bins * sh.angularBins : 0, 0);
	ls1g_rdf_params prm{};
	prm.bins = static_cast<uint32_t>(sh.bins);
	prm.angular_bins = sh.angular ? static_cast<uint32_t>(sh.angularBins) : 1;
	prm.interval_length = sh.interval;
	// the fused step has already dropped the halo copies (and may sit in the neighbour-list layout): put them back for this
	// traversal, as they are when the reference's plugin runs (Simulation.cpp:1037-1086)
	const bool stepDone = _mode == FUSED && _ranThisStep;
	if (stepDone) chk(ls1g_exchange(_ctx), "ls1g_exchange");
	chk(ls1g_rdf(_ctx, &prm, mol.data(), site.empty() ? nullptr : site.data(), ardf.empty() ? nullptr : ardf.data()), "ls1g_rdf");
	if (stepDone) chk(ls1g_delete_outer(_ctx), "ls1g_delete_outer");
	RDFTest::add(*rdf, comps, mol, site, ardf);
	++_rdfSamples;
	_tDevice += wall() - t0;
}

void GpuLinkedCells::traverseNonInnermostCells(CellProcessor&) {
	global_log->error_always_output() << "GpuLinkedCells: overlapping traversal is not supported (set overlappingP2P off)" << std::endl;
	Simulation::exit(1);
}

void GpuLinkedCells::traversePartialInnermostCells(CellProcessor&, unsigned int, int) {
	global_log->error_always_output() << "GpuLinkedCells: overlapping traversal is not supported (set overlappingP2P off)" << std::endl;
	Simulation::exit(1);
}

void GpuLinkedCells::deleteOuterParticles() {
	if (!_devicePhase) {
		LinkedCells::deleteOuterParticles();
		if (_deviceLoaded) chk(ls1g_delete_outer(_ctx), "ls1g_delete_outer");
		return;
	}
	if (_mode == OPERATORS || !_ranThisStep) {
		const double t0 = wall();
		chk(ls1g_delete_outer(_ctx), "ls1g_delete_outer");
		_tDevice += wall() - t0;
	}
}

unsigned long GpuLinkedCells::getNumberOfParticles() {
	if (!_devicePhase || _mirrorValid) return LinkedCells::getNumberOfParticles();
	uint64_t owned = 0, halo = 0;
	chk(ls1g_count(_ctx, &owned, &halo), "ls1g_count");
	return owned + halo;
}

void GpuLinkedCells::clear() {
	LinkedCells::clear();
	if (_ctx != nullptr && _deviceLoaded) {
		uint64_t id = 0;
		uint32_t cid = 0;
		double r[3] = {0., 0., 0.};
		chk(ls1g_upload(_ctx, 0, &id, &cid, r, nullptr, nullptr, nullptr), "ls1g_upload");
	}
	_deviceLoaded = false;
	_devicePhase = false;
	_mirrorValid = false;
}

bool GpuLinkedCells::mirrorWanted() const {
	const unsigned long step = global_simulation->getSimulationStep();
	if (step >= global_simulation->getNumTimesteps()) return true;       // final checkpoint / plugin finish
	return _mirrorFrequency > 0 && step % _mirrorFrequency == 0;
}

// Leapfrog::eventNewTimestep (Leapfrog.cpp:48-68)
void GpuLinkedCells::integratePre(double dt, Domain* domain) {
	_dt = dt;
	if (!_deviceLoaded) {
		global_log->error_always_output() << "GpuLinkedCells: time step without an initial force calculation" << std::endl;
		Simulation::exit(1);
	}
	if (!_devicePhase || _mirrorValid) {
		// from here on the device is authoritative: drop the host copies so that the host-side passes of the time loop
		// (exchangeMolecules, updateForces, thermostat apply) have nothing to do
		clearHostCells();
		_devicePhase = true;
		_mirrorValid = false;
	}
	const double t0 = wall();
	if (_steps > 0 && !domain->NVE()) {
		// velocity scaling of the previous step (Simulation.cpp:1108-1136) with the betas Domain::calculateGlobalValues
		// computed (Domain.cpp:226-247); nothing reads v in between
		const double bt = domain->getGlobalBetaTrans();
		const double br = domain->getGlobalBetaRot();
		// "Explosion heuristics" (Domain.cpp:254-311): when a raw beta is below MIN_BETA, calculateGlobalValues caps the kinetic
		// energy per molecule and replaces the betas by (1, beta_rot^0.0091).  Its molecule loop ran over the (empty) host
		// cells; the same capping is done here on the device.  The raw betas are recomputed from the very sums handed to Domain.
		const double Ti = _lastTfactor * domain->getTargetTemperature(0);
		if (Ti > 0.0 && _n > 0) {
			const double rawBt = std::pow(3.0 * _n * Ti / _summv2, 0.4);
			const double rawBr = (_sumIw2 == 0.0 || _rotDOF == 0) ? 1.0 : std::pow(_rotDOF * Ti / _sumIw2, 0.4);
			if ((rawBt < MIN_BETA || rawBr < MIN_BETA) && bt == 1.0 && std::fabs(br - std::pow(rawBr, 0.0091)) <= 1e-14) {
				chk(ls1g_selective_thermostat(_ctx, KINLIMIT_PER_T * Ti, Ti), "ls1g_selective_thermostat");
				++_selective;
			}
		}
		if (_mode == FUSED) {          // applied inside ls1g_step, fused into the pre-kick
			_pendingBt = bt;
			_pendingBr = br;
		} else if (bt != 1.0 || br != 1.0) {
			chk(ls1g_scale(_ctx, bt, br), "ls1g_scale");
		}
	}
	if (_mode == OPERATORS) chk(ls1g_integrate_pre(_ctx, dt), "ls1g_integrate_pre");
	_tDevice += wall() - t0;
	_inStep = true;
	_ranThisStep = false;
}

// Leapfrog::eventForcesCalculated (Leapfrog.cpp:70-155)
void GpuLinkedCells::integratePost(double dt, Domain* domain) {
	if (!_devicePhase) {
		global_log->error_always_output() << "GpuLinkedCells: eventForcesCalculated without eventNewTimestep" << std::endl;
		Simulation::exit(1);
	}
	const double t0 = wall();
	if (_mode == OPERATORS) {
		chk(ls1g_integrate_post(_ctx, 0.5 * dt, &_summv2, &_sumIw2, &_n, &_rotDOF), "ls1g_integrate_post");
	} else if (!_ranThisStep) {
		global_log->error_always_output() << "GpuLinkedCells (fused mode): time step without traverseCells" << std::endl;
		Simulation::exit(1);
	}
	_tDevice += wall() - t0;
	domain->setLocalSummv2(_summv2, 0);                  // Leapfrog.cpp:146-150
	domain->setLocalSumIw2(_sumIw2, 0);
	domain->setLocalNrotDOF(0, _n, _rotDOF);
	_inStep = false;
	_ranThisStep = false;
	_lastTfactor = global_simulation->Tfactor(global_simulation->getSimulationStep());   // the factor calculateGlobalValues gets next
	++_steps;
	if (mirrorWanted()) downloadMirror();
	if (global_simulation->getSimulationStep() >= global_simulation->getNumTimesteps()) printSummary();
}

// device -> host cells.  The download is cell-sorted on the reference grid, so threads fill disjoint cells.
void GpuLinkedCells::downloadMirror() {
	if (_ctx == nullptr || !_deviceLoaded) return;
	const double t0 = wall();
	clearHostCells();
	uint64_t slots = 0;
	chk(ls1g_slot_count(_ctx, &slots), "ls1g_slot_count");     // owned molecules + the slots of the (dropped) halo copies
	const size_t ns = static_cast<size_t>(slots);
	const double tStage = ensureStaging(ns + 1);
	uint64_t* id = _sId;
	uint32_t *cid = _sCid, *cell = _sCell;
	double *r = _sR, *v = _sV, *q = _sQ, *L = _sL, *F = _sF, *M = _sM, *Vi = _sVi;
	// a single LJ centre at the origin has no orientation on the device: q = (1, 0, 0, 0), D = M = 0 need not cross PCIe
	const bool point = _pointMolecules;
	const int mask = LS1G_F_ID | LS1G_F_CID | LS1G_F_R | LS1G_F_V | LS1G_F_F | LS1G_F_VI | (point ? 0 : (LS1G_F_Q | LS1G_F_L | LS1G_F_M));
	const double tCopy0 = wall();
	if (ns > 0) {	// the cell index of every slot first (4 bytes each): the chunks of the pipeline below are cut at cell borders
		ls1g_fields fc{};
		fc.cell = cell;
		chk(ls1g_download_range(_ctx, LS1G_F_CELL, 0, ns, &fc), "ls1g_download_range");
	}
	double tCopy = wall() - tCopy0;
	const double tFill0 = wall();
	// first slot of every run of equal cell index (halo runs are skipped when filling)
	std::vector<size_t> runs;
	{
		int nt = 1;
#if defined(_OPENMP)
		nt = omp_get_max_threads();
#endif
		std::vector<std::vector<size_t>> part(nt);
#if defined(_OPENMP)
#pragma omp parallel num_threads(nt)
#endif
		{
			int t = 0;
#if defined(_OPENMP)
			t = omp_get_thread_num();
#endif
			const size_t a = ns * t / nt, b = ns * (t + 1) / nt;
			for (size_t i = a; i < b; ++i)
				if (i == 0 || cell[i] != cell[i - 1]) part[t].push_back(i);
		}
		for (int t = 0; t < nt; ++t) runs.insert(runs.end(), part[t].begin(), part[t].end());
	}
	runs.push_back(ns);
	const long nruns = static_cast<long>(runs.size()) - 1;
	const long nchunks = std::max<long>(1, std::min<long>(16, static_cast<long>(ns / 262144)));
	const long rpc = std::max<long>(1, (nruns + nchunks - 1) / nchunks);      // runs per chunk
	const long block = 64;
	std::atomic<long> ready(0), cursor(0);
	std::vector<Component>& comps = *global_simulation->getEnsemble()->getComponents();
	const double zero3[3] = {0., 0., 0.};
	size_t n = 0;
	// thread 0 brings chunk after chunk over PCIe (ls1g_download_range blocks until the chunk is in host memory) while the other
	// threads refill the host cells of the chunks that have arrived; then it joins them
#if defined(_OPENMP)
#pragma omp parallel reduction(+ : n)
#endif
	{
		bool copier = true;
#if defined(_OPENMP)
		copier = omp_get_thread_num() == 0;
#endif
		if (copier) {
			const double tc = wall();
			for (long k = 0; k * rpc < nruns; ++k) {
				const long ra = k * rpc, rb = std::min((k + 1) * rpc, nruns);
				const size_t s0 = runs[ra], cnt = runs[rb] - s0;
				ls1g_fields f{};
				f.id = id + s0;
				f.cid = cid + s0;
				f.r = r + 3 * s0;
				f.v = v + 3 * s0;
				f.q = point ? nullptr : q + 4 * s0;
				f.L = point ? nullptr : L + 3 * s0;
				f.F = F + 3 * s0;
				f.M = point ? nullptr : M + 3 * s0;
				f.Vi = Vi + 3 * s0;
				chk(ls1g_download_range(_ctx, mask, s0, cnt, &f), "ls1g_download_range");
				ready.store(k + 1, std::memory_order_release);
			}
			tCopy += wall() - tc;
		}
		for (;;) {
			const long a = cursor.fetch_add(block);
			if (a >= nruns) break;
			const long b = std::min(a + block, nruns);
			const long need = (b - 1) / rpc + 1;
			while (ready.load(std::memory_order_acquire) < need) std::this_thread::yield();
			for (long k = a; k < b; ++k) {
				if ((cell[runs[k]] & LS1G_HALO_FLAG) != 0u) continue;
				ParticleCell& pc = getCellReference(static_cast<int>(cell[runs[k]]));
				pc.increaseMoleculeStorage(runs[k + 1] - runs[k]);
				for (size_t i = runs[k]; i < runs[k + 1]; ++i) {
					Molecule m = point ? Molecule(id[i], &comps[cid[i]], r[3 * i], r[3 * i + 1], r[3 * i + 2], v[3 * i], v[3 * i + 1], v[3 * i + 2],
					                              1., 0., 0., 0., 0., 0., 0.)
					                   : Molecule(id[i], &comps[cid[i]], r[3 * i], r[3 * i + 1], r[3 * i + 2], v[3 * i], v[3 * i + 1], v[3 * i + 2],
					                              q[4 * i], q[4 * i + 1], q[4 * i + 2], q[4 * i + 3], L[3 * i], L[3 * i + 1], L[3 * i + 2]);
					m.setF(&F[3 * i]);
					m.setM(const_cast<double*>(point ? zero3 : &M[3 * i]));
					m.setVi(&Vi[3 * i]);
					pc.addParticle(m);
				}
				n += runs[k + 1] - runs[k];
			}
		}
	}
	_tD2H += tCopy;                  // overlapped with the refill
	_tFill += wall() - tFill0;
	_bytesDown += ns * (8 + 4 + 4 + (point ? 12 : 22) * 8);
	(void)n;
	_mirrorValid = true;
	++_mirrors;
	_tMirror += wall() - t0 - tStage;
}
