/*
 * GpuLinkedCells.cpp — see GpuLinkedCells.h.  Every device operation is one call of the C ABI (include/ls1g.h); this file
 * only converts between the reference's host objects (Molecule, Component, Domain) and plain arrays.
 */
#include "GpuLinkedCells.h"
#include "RecordText.h"

#include "Domain.h"
#include "Simulation.h"
#include "ensemble/EnsembleBase.h"
#include "io/RDF.h"
#include "particleContainer/adapter/FlopCounter.h"
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

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
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
		for (size_t r = _ctxs.size(); r-- > 0;) ls1g_destroy(_ctxs[r]);
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
	                   << _selective << " time(s); " << _rdfSamples << " RDF and " << _flopSamples << " FLOP sample(s) on the device; first steps [s] " << _tFirstSteps[0] << " " << _tFirstSteps[1] << " "
	                   << _tFirstSteps[2] << " " << _tFirstSteps[3] << ", slowest step " << _tMaxStep
	                   << "; upload = gather " << _tGather << " + copy " << _tH2D << ", mirror = copy " << _tD2H << " + fill " << _tFill
	                   << "; " << _recordWrites << " record checkpoint(s) from the device took " << _tRecords << " s (" << _bytesRecords << " B)" << std::endl;
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
	std::string unknown = "refuse";
	xmlconfig.getNodeValue("unknownplugins", unknown);
	if (unknown != "refuse" && unknown != "readonly") {
		global_log->error_always_output() << "GpuLinkedCells: <unknownplugins> is refuse (default) or readonly, not " << unknown << std::endl;
		Simulation::exit(1);
	}
	_allowUnknownPlugins = unknown == "readonly";
	_mirrorFrequency = mirrorFrequencyFromPlugins(xmlconfig);      // also refuses plugins that are not known to be read-only
	if (mirror != "auto") {
		_mirrorFrequency = static_cast<unsigned long>(xmlconfig.getNodeValue_long("mirrorfrequency", 0));
	}
	_device = xmlconfig.getNodeValue_int("device", 0);
	_devices = xmlconfig.getNodeValue_int("devices", 1);
	if (_devices < 1) {
		global_log->error_always_output() << "GpuLinkedCells: <devices> must be >= 1" << std::endl;
		Simulation::exit(1);
	}
	global_log->info() << "GpuLinkedCells: mode " << mode << ", mirror frequency " << _mirrorFrequency
	                   << (mirror == "auto" ? " (from the plugins' write frequencies)" : "") << ", device " << _device << ", blocks " << _devices << std::endl;
}

// <mirrorfrequency>auto</mirrorfrequency> (the default): the end-of-step plugins that read molecules decide on which steps
// the host cells must hold a snapshot (PluginBase::endStep, Simulation.cpp:1150-1158).  Plugins that only use Domain
// values or timers are ignored; known molecule readers contribute their <writefrequency>; any other plugin is assumed to
// read molecules on every step.  Mirrors are read-only snapshots: a plugin that CHANGES molecules is not supported, and
// the ones that act before or after the force pass (PluginBase::beforeForces/afterForces) never see molecules.
unsigned long GpuLinkedCells::mirrorFrequencyFromPlugins(XMLfileUnits& xmlconfig) const {
	static const std::set<std::string> noMolecules = {"ResultWriter", "SysMonOutput", "TimerWriter", "LoadbalanceWriter",
	                                                  "DecompWriter", "CommunicationPartnerWriter", "KDTreePrinter",
	                                                  "RDF", "FlopRateWriter",       // their traversals are counted on the device (sampleRDF, sampleFlops)
	                                                  "GpuCheckpointWriter"};        // takes the packed records from the device (writeRecords)
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
			} else if (_allowUnknownPlugins) {
				global_log->warning() << "GpuLinkedCells: plugin " << name << " is not known to be a read-only end-of-step writer "
				                      << "(<unknownplugins>readonly</unknownplugins>): the host cells are refilled on every step for "
				                      << "it, changes it makes to molecules are NOT carried to the device" << std::endl;
			} else {
				// a plugin that moves, adds, deletes or rescales molecules would act on a snapshot and silently have no effect on
				// the trajectory: that is a wrong simulation, not a slow one, so it ends the run
				global_log->error_always_output() << "GpuLinkedCells: plugin " << name << " is not known to be a read-only end-of-step "
				                                  << "writer.  Plugins that change molecules are not supported by the device container "
				                                  << "(their changes would not reach the device).  If " << name << " only READS molecules, "
				                                  << "add <unknownplugins>readonly</unknownplugins> to the <datastructure> element."
				                                  << std::endl;
				Simulation::exit(1);
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
		for (size_t r = _ctxs.size(); r-- > 0;) ls1g_destroy(_ctxs[r]);
		_ctxs.clear();
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
	// block grid of the box: one block (the sequential DomainDecompBase) or DomainDecomposition's grid over <devices> blocks
	int ndev = 0;
	chk(ls1g_device_count(&ndev), "ls1g_device_count");
	if (ndev < 1) fail("no CUDA device");
	double L[3];
	for (int d = 0; d < 3; ++d) L[d] = domain->getGlobalLength(d);
	_ctxs.assign(static_cast<size_t>(_devices), nullptr);
	for (int r = 0; r < _devices; ++r) {
		ls1g_config cfg{};
		int grid[3], coords[3];
		chk(ls1g_decomp_grid(_devices, r, grid, coords), "ls1g_decomp_grid");
		cfg.device = (_device + r) % ndev;
		for (int d = 0; d < 3; ++d) {
			cfg.global_length[d] = L[d];
			cfg.rank_grid[d] = grid[d];
			cfg.rank_coords[d] = coords[d];
			_blockGrid[d] = grid[d];
		}
		if (_devices == 1) {
			for (int d = 0; d < 3; ++d) {
				cfg.box_min[d] = getBoundingBoxMin(d);
				cfg.box_max[d] = getBoundingBoxMax(d);
			}
		} else {
			chk(ls1g_decomp_box(grid, coords, L, cfg.box_min, cfg.box_max), "ls1g_decomp_box");
		}
		cfg.cutoff = getCutoff();
		cfg.lj_cutoff = global_simulation->getLJCutoff();
		cfg.eps_rf = domain->getepsilonRF();
		cfg.grid_mode = LS1G_GRID_REBUILD;
		cfg.capacity = 0;
		chk(ls1g_create(&cfg, &_ctxs[r]), "ls1g_create");
	}
	_ctx = _ctxs[0];
	if (multi()) {
		for (int d = 0; d < 3; ++d) {      // block bounds per dimension, exactly as the contexts got them
			_blockLo[d].assign(static_cast<size_t>(_blockGrid[d]) + 1, 0.);
			for (int k = 0; k < _blockGrid[d]; ++k) {
				int grid[3] = {_blockGrid[0], _blockGrid[1], _blockGrid[2]}, co[3] = {0, 0, 0};
				double lo[3], hi[3];
				co[d] = k;
				chk(ls1g_decomp_box(grid, co, L, lo, hi), "ls1g_decomp_box");
				_blockLo[d][k] = lo[d];
				_blockLo[d][k + 1] = hi[d];
			}
		}
	}
	uploadComponents();
	if (multi()) chk(ls1g_comm_init_local(_ctxs.data(), _devices), "ls1g_comm_init_local");
	int cpd[3];
	double len[3];
	chk(ls1g_get_grid(_ctx, cpd, len), "ls1g_get_grid");
	const double* mine = getCellLength();
	for (int d = 0; d < 3 && !multi(); ++d) {
		if (len[d] != mine[d]) {
			global_log->error_always_output() << "GpuLinkedCells: device cell grid differs from LinkedCells::rebuild in dimension "
			                                  << d << " (" << len[d] << " vs " << mine[d] << ")" << std::endl;
			Simulation::exit(1);
		}
	}
	if (multi()) {      // the host cells keep the whole box: their grid is LinkedCells::rebuild's, the blocks have their own
		const int* w = getBoxWidthInNumCells();
		for (int d = 0; d < 3; ++d) cpd[d] = w[d] + 2;
	}
	_numCells = static_cast<size_t>(cpd[0]) * cpd[1] * cpd[2];
	global_log->info() << "GpuLinkedCells: " << _devices << " device context(s) created (block grid " << _blockGrid[0] << " x " << _blockGrid[1] << " x "
	                   << _blockGrid[2] << " on " << ndev << " visible GPU(s)), " << cpd[0] << " x " << cpd[1] << " x " << cpd[2] << " host cells" << std::endl;
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
	for (ls1g_ctx* cx : _ctxs)
		chk(ls1g_set_components(cx, static_cast<int>(nc), rows.data(), mix.empty() ? nullptr : mix.data(),
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
	if (!_deviceLoaded && !multi()) {             // ... and so is the allocation of the device arrays
		const double tr = wall();
		chk(ls1g_reserve(_ctx, n), "ls1g_reserve");
		_tContext += wall() - tr;
		tStage += wall() - tr;
	}
	uint64_t* id = _sId;
	uint32_t* cid = _sCid;
	double *r = _sR, *v = _sV, *q = _sQ, *L = _sL;
	const bool point = _pointMolecules;      // the device keeps no orientation for a single LJ centre at the origin
	auto put = [=](const Molecule& m, size_t i) {
		id[i] = m.getID();
		cid[i] = m.componentid();
		for (int d = 0; d < 3; ++d) {
			r[3 * i + d] = m.r(d);
			v[3 * i + d] = m.v(d);
		}
		if (point) return;
		for (int d = 0; d < 3; ++d) L[3 * i + d] = m.D(d);
		const Quaternion& qq = m.q();
		q[4 * i] = qq.qw();
		q[4 * i + 1] = qq.qx();
		q[4 * i + 2] = qq.qy();
		q[4 * i + 3] = qq.qz();
	};
	std::vector<size_t> blockStart;
	if (!multi()) {
#if defined(_OPENMP)
#pragma omp parallel for schedule(static)
#endif
		for (long c = 0; c < nc; ++c) {
			if (first[c + 1] == first[c]) continue;
			size_t i = first[c];
			ParticleCell& pc = getCellReference(static_cast<int>(c));
			for (auto m = pc.iterator(); m.isValid(); ++m, ++i) put(*m, i);
		}
	} else {
		// <devices> > 1: the staging arrays are filled block by block (a counting sort by owning block, all threads: every thread
		// counts the molecules of its cells per block, the prefix sums give it a private range in every block, then it writes)
		const int nb = numDevices();
		int nt = 1;
#if defined(_OPENMP)
		nt = omp_get_max_threads();
#endif
		std::vector<size_t> cursor(static_cast<size_t>(nt) * nb, 0);
		std::vector<int> cellBlock(static_cast<size_t>(nc), -1);
		blockStart.assign(static_cast<size_t>(nb) + 1, 0);
#if defined(_OPENMP)
#pragma omp parallel num_threads(nt)
#endif
		{
			int t = 0;
#if defined(_OPENMP)
			t = omp_get_thread_num();
#endif
			const long c0 = nc * t / nt, c1 = nc * (t + 1) / nt;
			size_t* mine = cursor.data() + static_cast<size_t>(t) * nb;
			// a host cell that lies inside one block (all of them when the block faces fall on cell faces) is counted without
			// touching its molecules; only a cell cut by a block face is walked molecule by molecule
			for (long c = c0; c < c1; ++c) {
				if (first[c + 1] == first[c]) continue;
				ParticleCell& pc = getCellReference(static_cast<int>(c));
				const int whole = blockOfCell(pc);
				cellBlock[c] = whole;
				if (whole >= 0) { mine[whole] += first[c + 1] - first[c]; continue; }
				for (auto m = pc.iterator(); m.isValid(); ++m) ++mine[blockOf(m->r_arr().data())];
			}
#if defined(_OPENMP)
#pragma omp barrier
#pragma omp single
#endif
			{
				for (int b = 0; b < nb; ++b) {
					size_t run = blockStart[b];
					for (int u = 0; u < nt; ++u) {
						size_t& cu = cursor[static_cast<size_t>(u) * nb + b];
						const size_t cnt = cu;
						cu = run;
						run += cnt;
					}
					blockStart[b + 1] = run;
				}
			}
			for (long c = c0; c < c1; ++c) {
				if (first[c + 1] == first[c]) continue;
				ParticleCell& pc = getCellReference(static_cast<int>(c));
				const int whole = cellBlock[c];
				if (whole >= 0) { for (auto m = pc.iterator(); m.isValid(); ++m) put(*m, mine[whole]++); }
				else { for (auto m = pc.iterator(); m.isValid(); ++m) put(*m, mine[blockOf(m->r_arr().data())]++); }
			}
		}
	}
	const double tCopy0 = wall();
	_tGather += tCopy0 - t0 - tStage;
	if (multi()) uploadBlocks(blockStart);
	else chk(ls1g_upload(_ctx, n, id, cid, r, v, _pointMolecules ? nullptr : q, _pointMolecules ? nullptr : L), "ls1g_upload");
	_tH2D += wall() - tCopy0;
	_bytesUp += n * (8 + 4 + (_pointMolecules ? 6 : 13) * 8);
	_deviceLoaded = true;
	_tUpload += wall() - t0 - tStage;
}

// block (rank, row-major over the block grid as MPI_Cart_rank) that owns the point r: [lo, hi) per dimension with the bounds the
// contexts were created with
int GpuLinkedCells::blockOf(const double* r) const {
	int co[3];
	for (int d = 0; d < 3; ++d) {
		const std::vector<double>& lo = _blockLo[d];
		const int g = _blockGrid[d];
		int k = static_cast<int>(r[d] / (lo[g] - lo[0]) * g);
		k = std::max(0, std::min(g - 1, k));
		while (k > 0 && r[d] < lo[k]) --k;
		while (k + 1 < g && r[d] >= lo[k + 1]) ++k;
		co[d] = k;
	}
	return (co[0] * _blockGrid[1] + co[1]) * _blockGrid[2] + co[2];
}

// the block that holds the whole host cell, or -1 when a block face cuts through it (its molecules are then assigned one by one).
// A molecule exactly on the cell's upper face belongs to the next cell, so [min, max) inside [lo, hi) is the test.
int GpuLinkedCells::blockOfCell(const ParticleCell& pc) const {
	int co[3];
	for (int d = 0; d < 3; ++d) {
		const std::vector<double>& lo = _blockLo[d];
		const int g = _blockGrid[d];
		const double a = pc.getBoxMin(d), b = pc.getBoxMax(d);
		int k = 0;
		while (k + 1 < g && a >= lo[k + 1]) ++k;
		if (k + 1 < g && b > lo[k + 1]) return -1;
		co[d] = k;
	}
	return (co[0] * _blockGrid[1] + co[1]) * _blockGrid[2] + co[2];
}

// <devices> > 1: the staging arrays hold the molecules block by block (start[b] .. start[b + 1]); one upload per block, straight from
// the page-locked staging
void GpuLinkedCells::uploadBlocks(const std::vector<size_t>& start) {
	const int nb = numDevices();
	const bool rot = !_pointMolecules;
	// every block is given room for the fullest one: its halo copies come from neighbours that may hold more molecules than itself
	size_t fullest = 0;
	for (int b = 0; b < nb; ++b) fullest = std::max(fullest, start[b + 1] - start[b]);
	if (!_deviceLoaded) {
		const double tr = wall();
		for (int b = 0; b < nb; ++b) chk(ls1g_reserve(_ctxs[b], fullest + fullest / 8), "ls1g_reserve");
		_tContext += wall() - tr;
	}
	// one host thread per block: the copies to the GPUs run side by side (each GPU has its own PCIe link)
	std::vector<int> failed(static_cast<size_t>(nb), 0);
#if defined(_OPENMP)
#pragma omp parallel for schedule(static, 1) num_threads(nb)
#endif
	for (int b = 0; b < nb; ++b) {
		const size_t m = start[b + 1] - start[b], o = start[b];
		uint64_t id0 = 0; uint32_t cid0 = 0; double r0[3] = {0., 0., 0.};
		if (ls1g_upload(_ctxs[b], m, m ? _sId + o : &id0, m ? _sCid + o : &cid0, m ? _sR + 3 * o : r0, m ? _sV + 3 * o : nullptr,
		                rot && m ? _sQ + 4 * o : nullptr, rot && m ? _sL + 3 * o : nullptr) != 0) failed[b] = 1;
	}
	for (int b = 0; b < nb; ++b) if (failed[b]) fail("ls1g_upload");
}

void GpuLinkedCells::exchangeAll() {
	if (multi()) chk(ls1g_exchange_group(_ctxs.data(), numDevices()), "ls1g_exchange_group");
	else chk(ls1g_exchange(_ctx), "ls1g_exchange");
}

void GpuLinkedCells::traverseAll(double& upot, double& virial) {
	upot = virial = 0.;
	for (ls1g_ctx* cx : _ctxs) {
		double u = 0., w = 0.;
		chk(ls1g_traverse(cx, &u, &w), "ls1g_traverse");
		upot += u;
		virial += w;
	}
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
		exchangeAll();
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
		if (_devicePhase && !_mirrorValid) {
			if (FlopCounter* fc = dynamic_cast<FlopCounter*>(&cellProcessor)) {
				sampleFlops(*fc);      // FlopRateWriter::afterForces: the reference's FLOP model over device-side counts
				return;
			}
		}
		// other counting / sampling processors run on the host cells
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
		exchangeAll();
		traverseAll(upot, virial);
		_tDevice += wall() - t0;
	} else if (_mode == FUSED && _inStep && !_ranThisStep) {
		const double t0 = wall();
		ls1g_step_record rec{};
		// [thermostat of the previous step] -> pre -> exchange -> forces (+calcFM) -> deleteOuter -> post, as one device-resident
		// step; the betas are the Domain's own (integratePre).  Vi and the pair count only when the host will read them.
		if (multi()) chk(ls1g_step_group(_ctxs.data(), numDevices(), _dt, _pendingBt, _pendingBr, mirrorWanted() ? LS1G_STEP_VI : 0, &rec), "ls1g_step_group");
		else chk(ls1g_step(_ctx, _dt, _pendingBt, _pendingBr, mirrorWanted() ? LS1G_STEP_VI : 0, &rec), "ls1g_step");
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
		traverseAll(upot, virial);
		_tDevice += wall() - t0;
	}
	domain->setLocalUpot(upot);          // VectorizedCellProcessor::endTraversal (VectorizedCellProcessor.cpp:155-156)
	domain->setLocalVirial(virial);
}

// RDF::deviceShape / RDF::addLocalCounts: the two public methods a maintainer adds to io/RDF.h for containers that count the pair
// histograms elsewhere (here the declaration is inserted into a scratch copy of the header by
// host/Makefile, this is the definition).  One traversal's counts go to the plugin's LOCAL counters, which RDF::collectRDF then
// sums over ranks and RDF::accumulateRDF adds to the running total (io/RDF.cpp:249-330) exactly as after its own cell processor.
//   mol  [pair row (i, i+k)][bin]   site [global site m][global site n][bin]   ardf [i][k][bin * angular bin]
void RDF::addLocalCounts(const std::vector<Component>& comps, const std::vector<uint64_t>& mol, const std::vector<uint64_t>& site,
                         const std::vector<uint64_t>& ardf) {
	const size_t nC = comps.size(), bins = _bins;
	std::vector<size_t> site0(nC + 1, 0);
	for (size_t i = 0; i < nC; ++i) site0[i + 1] = site0[i] + comps[i].numSites();
	const size_t S = site0[nC];
	size_t row = 0;
	for (size_t i = 0; i < nC; ++i) {
		for (size_t k = 0; i + k < nC; ++k, ++row) {
			for (size_t l = 0; l < bins; ++l) _distribution.local[i][k][l] += mol[row * bins + l];
			const size_t ni = comps[i].numSites(), nj = comps[i + k].numSites();
			if (ni + nj > 2) {
				for (size_t m = 0; m < ni; ++m)
					for (size_t n = 0; n < nj; ++n)
						for (size_t l = 0; l < bins; ++l)
							_siteDistribution.local[i][k][m][n][l] += site[((site0[i] + m) * S + site0[i + k] + n) * bins + l];
			}
		}
		if (_doARDF) {
			const size_t per = bins * _angularBins;
			for (size_t k = 0; k < nC; ++k)
				for (size_t l = 0; l < per; ++l) _ARDFdistribution.local[i][k][l] += ardf[(i * nC + k) * per + l];
		}
	}
}

// FlopRateWriter::measureFLOPS (io/FlopRateWriter.cpp:130-138) hands its FlopCounter to traverseCells after the force pass.  The
// device counts the same quantities over the same cell pairs (ls1g_flop_count, flops.cu); they enter the counter through
// FlopCounter::addDeviceCounts (the second upstream addition, see host/Makefile) between its own initTraversal / endTraversal, so
// the FLOP model, the sum over ranks and the writer's output stay the reference's code.
void GpuLinkedCells::sampleFlops(FlopCounter& counter) {
	const double t0 = wall();
	const bool stepDone = _mode == FUSED && _ranThisStep;      // halo copies dropped by the fused step: put them back (see sampleRDF)
	if (stepDone) exchangeAll();
	counter.initTraversal();
	for (ls1g_ctx* cx : _ctxs) {
		ls1g_flop_counts fc{};
		chk(ls1g_flop_count(cx, &fc), "ls1g_flop_count");
		counter.addDeviceCounts(fc.molecule_distances, fc.kernel_calls, fc.macro_calls);
	}
	counter.endTraversal();
	if (stepDone) for (ls1g_ctx* cx : _ctxs) chk(ls1g_delete_outer(cx), "ls1g_delete_outer");
	++_flopSamples;
	_tDevice += wall() - t0;
}

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
	const RDF::DeviceShape sh = rdf->deviceShape();
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
	if (stepDone) exchangeAll();
	for (ls1g_ctx* cx : _ctxs)       // every block adds the pairs the reference counts on that rank (RDF::collectRDF sums over ranks)
		chk(ls1g_rdf(cx, &prm, mol.data(), site.empty() ? nullptr : site.data(), ardf.empty() ? nullptr : ardf.data()), "ls1g_rdf");
	if (stepDone) for (ls1g_ctx* cx : _ctxs) chk(ls1g_delete_outer(cx), "ls1g_delete_outer");
	rdf->addLocalCounts(comps, mol, site, ardf);
	++_rdfSamples;
	_tDevice += wall() - t0;
}

// The reference's overlap API (C08CellPairTraversal.h:78-199: traverseCellPairsOuter / traverseCellPairsInner, driven by
// NonBlockingMPIMultiStepHandler through Simulation::performOverlappingDecompositionAndCellTraversalStep, Simulation.cpp:1301-1319 -
// a path that exists with ENABLE_MPI only).  Its purpose - compute the inner cells while the halo travels - is what the device step
// does by itself: the force kernel runs as phase 1 (blocks whose tile touches no halo cell) while the halo refresh is in flight on
// a second stream, then phase 2 (boundary blocks) (lists.cuh, step.cu).  So whichever of the two entry points the handler
// calls first in a time step runs that step; the other calls of the step find the work done.
void GpuLinkedCells::traverseNonInnermostCells(CellProcessor& cellProcessor) {
	if (_devicePhase && _mode == FUSED && _inStep && _ranThisStep) return;
	traverseCells(cellProcessor);
}

void GpuLinkedCells::traversePartialInnermostCells(CellProcessor& cellProcessor, unsigned int, int) {
	if (_devicePhase && _mode == FUSED && _inStep && _ranThisStep) return;
	traverseCells(cellProcessor);
}

void GpuLinkedCells::deleteOuterParticles() {
	if (!_devicePhase) {
		LinkedCells::deleteOuterParticles();
		if (_deviceLoaded) for (ls1g_ctx* cx : _ctxs) chk(ls1g_delete_outer(cx), "ls1g_delete_outer");
		return;
	}
	if (_mode == OPERATORS || !_ranThisStep) {
		const double t0 = wall();
		for (ls1g_ctx* cx : _ctxs) chk(ls1g_delete_outer(cx), "ls1g_delete_outer");
		_tDevice += wall() - t0;
	}
}

unsigned long GpuLinkedCells::getNumberOfParticles() {
	if (!_devicePhase || _mirrorValid) return LinkedCells::getNumberOfParticles();
	uint64_t total = 0;
	for (ls1g_ctx* cx : _ctxs) {
		uint64_t owned = 0, halo = 0;
		chk(ls1g_count(cx, &owned, &halo), "ls1g_count");
		total += owned + halo;
	}
	return total;
}

void GpuLinkedCells::clear() {
	LinkedCells::clear();
	if (_ctx != nullptr && _deviceLoaded) {
		uint64_t id = 0;
		uint32_t cid = 0;
		double r[3] = {0., 0., 0.};
		for (ls1g_ctx* cx : _ctxs) chk(ls1g_upload(cx, 0, &id, &cid, r, nullptr, nullptr, nullptr), "ls1g_upload");
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
		applySelectiveThermostat(domain);
		if (_mode == FUSED) {          // applied inside ls1g_step, fused into the pre-kick
			_pendingBt = bt;
			_pendingBr = br;
		} else if (bt != 1.0 || br != 1.0) {
			for (ls1g_ctx* cx : _ctxs) chk(ls1g_scale(cx, bt, br), "ls1g_scale");
		}
	}
	if (_mode == OPERATORS) for (ls1g_ctx* cx : _ctxs) chk(ls1g_integrate_pre(cx, dt), "ls1g_integrate_pre");
	_tDevice += wall() - t0;
	_inStep = true;
	_ranThisStep = false;
}

// "Explosion heuristics" (Domain.cpp:254-311): when a raw beta is below MIN_BETA, calculateGlobalValues caps the kinetic energy per
// molecule and replaces the betas by (1, beta_rot^0.0091).  Its molecule loop ran over the (empty) host cells; the same capping is
// done here on the device.  The raw betas are recomputed from the very sums handed to Domain.  Runs once per time step: either when a
// checkpoint is taken from the device at the end of the step (writeRecords) or at the start of the next step (integratePre) -
// nothing touches v, D in between.
void GpuLinkedCells::applySelectiveThermostat(Domain* domain) {
	if (_selectiveStep == _steps) return;
	_selectiveStep = _steps;
	const double bt = domain->getGlobalBetaTrans();
	const double br = domain->getGlobalBetaRot();
	const double Ti = _lastTfactor * domain->getTargetTemperature(0);
	if (Ti > 0.0 && _n > 0) {
		const double rawBt = std::pow(3.0 * _n * Ti / _summv2, 0.4);
		const double rawBr = (_sumIw2 == 0.0 || _rotDOF == 0) ? 1.0 : std::pow(_rotDOF * Ti / _sumIw2, 0.4);
		if ((rawBt < MIN_BETA || rawBr < MIN_BETA) && bt == 1.0 && std::fabs(br - std::pow(rawBr, 0.0091)) <= 1e-14) {
			for (ls1g_ctx* cx : _ctxs) chk(ls1g_selective_thermostat(cx, KINLIMIT_PER_T * Ti, Ti), "ls1g_selective_thermostat");
			++_selective;
		}
	}
}

uint64_t GpuLinkedCells::writeRecords(const std::string& file, Domain* domain, bool ascii) {
	if (!deviceHoldsState()) {
		global_log->error_always_output() << "GpuLinkedCells::writeRecords: the host cells hold the state" << std::endl;
		Simulation::exit(1);
	}
	const double t0 = wall();
	// the reference's writer runs after VelocityScalingThermostat::apply (Simulation.cpp:1108-1136, plugins' endStep :1150-1158);
	// the device applies that scaling inside its next step, so it goes into the records here (same multiplication, same bits)
	double bt = 1.0, br = 1.0;
	if (_steps > 0 && !domain->NVE()) {
		applySelectiveThermostat(domain);
		bt = domain->getGlobalBetaTrans();
		br = domain->getGlobalBetaRot();
	}
	std::FILE* f = std::fopen(file.c_str(), ascii ? "ab" : "wb");
	if (f == nullptr) {
		global_log->error_always_output() << "GpuLinkedCells::writeRecords: could not open " << file << std::endl;
		Simulation::exit(1);
	}
	auto put = [&](const void* p, size_t bytes) {
		if (bytes != 0 && std::fwrite(p, 1, bytes, f) != bytes) {
			global_log->error_always_output() << "GpuLinkedCells::writeRecords: short write to " << file << std::endl;
			Simulation::exit(1);
		}
	};
	uint64_t total = 0;
	size_t written = 0;
	for (ls1g_ctx* cx : _ctxs) {
		size_t n = 0;
		chk(ls1g_download_records(cx, 1.0, 1.0, 0, nullptr, &n), "ls1g_download_records");
		if (n == 0) continue;
		ensureStaging(n + n / 28 + 16);                // the staging block holds >= 112 bytes per molecule: room for 116-byte records
		chk(ls1g_download_records(cx, bt, br, n, _stage, &n), "ls1g_download_records");
		const unsigned char* rec = static_cast<const unsigned char*>(_stage);
		if (!ascii) {
			put(rec, n * ls1g_records::kRecordBytes);
			written += n * ls1g_records::kRecordBytes;
		} else {
			// text in slabs of 2^20 molecules: every thread prints a contiguous share of the slab, the shares go to the file in order
			int nt = 1;
#if defined(_OPENMP)
			nt = omp_get_max_threads();
#endif
			std::vector<std::string> part(nt);
			const size_t slab = size_t(1) << 20;
			for (size_t first = 0; first < n; first += slab) {
				const size_t count = std::min(slab, n - first);
#if defined(_OPENMP)
#pragma omp parallel for schedule(static, 1)
#endif
				for (int t = 0; t < nt; ++t) {
					part[t].clear();
					ls1g_records::appendLines(rec, first + count * t / nt, first + count * (t + 1) / nt, part[t]);
				}
				for (int t = 0; t < nt; ++t) {
					put(part[t].data(), part[t].size());
					written += part[t].size();
				}
			}
		}
		total += n;
	}
	std::fclose(f);
	++_recordWrites;
	_bytesRecords += written;
	_tRecords += wall() - t0;
	return total;
}

// Leapfrog::eventForcesCalculated (Leapfrog.cpp:70-155)
void GpuLinkedCells::integratePost(double dt, Domain* domain) {
	if (!_devicePhase) {
		global_log->error_always_output() << "GpuLinkedCells: eventForcesCalculated without eventNewTimestep" << std::endl;
		Simulation::exit(1);
	}
	const double t0 = wall();
	if (_mode == OPERATORS) {
		_summv2 = _sumIw2 = 0.;
		_n = _rotDOF = 0;
		for (ls1g_ctx* cx : _ctxs) {
			double a = 0., b = 0.;
			uint64_t nn = 0, rd = 0;
			chk(ls1g_integrate_post(cx, 0.5 * dt, &a, &b, &nn, &rd), "ls1g_integrate_post");
			_summv2 += a; _sumIw2 += b; _n += nn; _rotDOF += rd;
		}
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
	if (multi()) { downloadMirrorBlocks(); return; }
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


// <devices> > 1: device blocks -> host cells.  The blocks have their own cell grids, so every owned molecule is put into the host
// cell LinkedCells::getCellIndexOfPoint gives for it: cell index per molecule, counting sort by host cell, then the threads fill
// disjoint cells (molecules of a cell in id order, so the host state does not depend on the thread schedule).
void GpuLinkedCells::downloadMirrorBlocks() {
	const double t0 = wall();
	clearHostCells();
	const bool point = _pointMolecules;
	const int mask = LS1G_F_ID | LS1G_F_CID | LS1G_F_R | LS1G_F_V | LS1G_F_F | LS1G_F_VI | (point ? 0 : (LS1G_F_Q | LS1G_F_L | LS1G_F_M));
	std::vector<Component>& comps = *global_simulation->getEnsemble()->getComponents();
	const double zero3[3] = {0., 0., 0.};
	const int nb = numDevices();
	// owned molecules per block (the count query also takes the group out of the list layout: one exchange for all blocks)
	std::vector<size_t> start(static_cast<size_t>(nb) + 1, 0);
	for (int b = 0; b < nb; ++b) {
		ls1g_fields none{};
		size_t m = 0;
		chk(ls1g_download(_ctxs[b], 0, 0, 0, &none, &m), "ls1g_download (count)");
		start[b + 1] = start[b] + m;
	}
	const size_t n = start[nb];
	const double tStageAll = ensureStaging(n + 1);
	// every block copies its molecules into its own range of the page-locked staging; one host thread per block, so the copies of
	// the GPUs run side by side (each GPU has its own PCIe link)
	const double tc = wall();
	std::vector<int> failed(static_cast<size_t>(nb), 0);
#if defined(_OPENMP)
#pragma omp parallel for schedule(static, 1) num_threads(nb)
#endif
	for (int b = 0; b < nb; ++b) {
		const size_t o = start[b], m = start[b + 1] - start[b];
		if (m == 0) continue;
		ls1g_fields f{};
		f.id = _sId + o; f.cid = _sCid + o; f.r = _sR + 3 * o; f.v = _sV + 3 * o; f.F = _sF + 3 * o; f.Vi = _sVi + 3 * o;
		if (!point) { f.q = _sQ + 4 * o; f.L = _sL + 3 * o; f.M = _sM + 3 * o; }
		size_t got = 0;
		if (ls1g_download(_ctxs[b], mask, 0, m, &f, &got) != 0 || got != m) failed[b] = 1;
	}
	for (int b = 0; b < nb; ++b) if (failed[b]) fail("ls1g_download");
	_tD2H += wall() - tc;
	_bytesDown += n * (8 + 4 + (point ? 12 : 22) * 8);
	const double tf = wall();
	// counting sort of all molecules by host cell, all threads (the order inside a cell is fixed by the id sort below)
	std::vector<uint32_t> hc(n), first(_numCells + 1, 0), fill(_numCells, 0), ord(n);
#if defined(_OPENMP)
#pragma omp parallel for schedule(static)
#endif
	for (long i = 0; i < static_cast<long>(n); ++i) {
		hc[i] = static_cast<uint32_t>(getCellIndexOfPoint(_sR + 3 * i));
#if defined(_OPENMP)
#pragma omp atomic
#endif
		++first[hc[i] + 1];
	}
	for (size_t c = 0; c < _numCells; ++c) first[c + 1] += first[c];
#if defined(_OPENMP)
#pragma omp parallel for schedule(static)
#endif
	for (long i = 0; i < static_cast<long>(n); ++i) {
		uint32_t k;
#if defined(_OPENMP)
#pragma omp atomic capture
#endif
		k = fill[hc[i]]++;
		ord[first[hc[i]] + k] = static_cast<uint32_t>(i);
	}
#if defined(_OPENMP)
#pragma omp parallel for schedule(dynamic, 256)
#endif
	for (long c = 0; c < static_cast<long>(_numCells); ++c) {
		const size_t a = first[c], b = first[c + 1];
		if (a == b) continue;
		std::sort(ord.begin() + a, ord.begin() + b, [&](uint32_t x, uint32_t y) { return _sId[x] < _sId[y]; });
		ParticleCell& pc = getCellReference(static_cast<int>(c));
		pc.increaseMoleculeStorage(b - a);
		for (size_t k = a; k < b; ++k) {
			const size_t i = ord[k];
			Molecule m = point ? Molecule(_sId[i], &comps[_sCid[i]], _sR[3 * i], _sR[3 * i + 1], _sR[3 * i + 2], _sV[3 * i], _sV[3 * i + 1], _sV[3 * i + 2],
			                              1., 0., 0., 0., 0., 0., 0.)
			                   : Molecule(_sId[i], &comps[_sCid[i]], _sR[3 * i], _sR[3 * i + 1], _sR[3 * i + 2], _sV[3 * i], _sV[3 * i + 1], _sV[3 * i + 2],
			                              _sQ[4 * i], _sQ[4 * i + 1], _sQ[4 * i + 2], _sQ[4 * i + 3], _sL[3 * i], _sL[3 * i + 1], _sL[3 * i + 2]);
			m.setF(&_sF[3 * i]);
			m.setM(const_cast<double*>(point ? zero3 : &_sM[3 * i]));
			m.setVi(&_sVi[3 * i]);
			pc.addParticle(m);
		}
	}
	_tFill += wall() - tf;
	_mirrorValid = true;
	++_mirrors;
	_tMirror += wall() - t0 - tStageAll;
}
