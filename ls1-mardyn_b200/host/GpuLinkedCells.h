/*
 * GpuLinkedCells — ls1-MarDyn ParticleContainer whose short-range force pass, re-sort, halo images and integration run on
 * one B200 through the C ABI in include/ls1g.h (libls1g.so).  Selected like every other container:
 *
 *   <datastructure type="GpuLinkedCells">
 *     <cellsInCutoffRadius>1</cellsInCutoffRadius>
 *     <mode>fused</mode>                    fused (default): one ls1g_run step per traverseCells; operators: one C-ABI call per reference operator
 *     <mirrorfrequency>auto</mirrorfrequency>  every k-th step the host cells are refilled from the device (writers, plugins); 0 = last step only;
 *                                              auto (default) = gcd of the write frequencies of the plugins that read molecules
 *     <device>0</device>
 *     <devices>1</devices>                  N > 1: the box is cut into DomainDecomposition's block grid (MPI_Dims_create order,
 *                                           DomainDecomposition.cpp:19-41,116-125) and block r lives on device (device + r) mod visible
 *                                           GPUs, all inside THIS process; halo and leaving molecules move between the blocks over
 *                                           NVLink peer copies (ls1g_comm_init_local, ls1g_step_group) instead of MPI
 *   </datastructure>
 *   <integrator type="GpuLeapfrog"> ... </integrator>
 *
 * This is host code of the reference's own kind: it is compiled against the reference headers (src/particleContainer/
 * LinkedCells.h, src/integrators/Integrator.h) and linked with the reference objects, see ls1-mardyn_b200/host/Makefile.
 * It derives from LinkedCells because ParticleIterator is a concrete class over std::vector<ParticleCell>
 * (src/particleContainer/ParticleIterator.h:50-90): readers, generators, writers and plugins keep working on the host
 * cells, which hold the molecules before the first time step ("host phase") and on mirror steps.  Precedent for a
 * container that ignores the cell processor's kernels and publishes upot/virial itself: AutoPasContainer::traverseCells
 * (src/particleContainer/AutoPasContainer.cpp:282-395).
 *
 * The unmodified time loop Simulation::simulate (src/Simulation.cpp:979-1177) drives it:
 *   eventNewTimestep -> update -> balanceAndExchange -> updateMoleculeCaches -> traverseCells -> updateForces ->
 *   deleteOuterParticles -> eventForcesCalculated -> calculateGlobalValues -> thermostat
 * In the device phase the host cells are empty, so the host-side passes of the loop (DomainDecompBase::exchangeMolecules,
 * Simulation::updateForces, VelocityScalingThermostat::apply) find nothing to do; their work happens in the calls above.
 */
#ifndef GPULINKEDCELLS_H_
#define GPULINKEDCELLS_H_

#include "particleContainer/LinkedCells.h"

#include <cstdint>
#include <string>
#include <vector>

struct ls1g_ctx;
class Domain;
class FlopCounter;

class GpuLinkedCells : public LinkedCells {
public:
	GpuLinkedCells();
	~GpuLinkedCells() override;

	void readXML(XMLfileUnits& xmlconfig) override;                         // LinkedCells.cpp:128-134
	bool rebuild(double bBoxMin[3], double bBoxMax[3]) override;            // LinkedCells.cpp:136-204
	void update() override;                                                 // LinkedCells.cpp:243-356
	void updateMoleculeCaches() override;                                   // LinkedCells.cpp:1076-1086
	void updateInnerMoleculeCaches() override { updateMoleculeCaches(); }
	void updateBoundaryAndHaloMoleculeCaches() override {}
	void traverseCells(CellProcessor& cellProcessor) override;              // LinkedCells.cpp:564-575
	void traverseNonInnermostCells(CellProcessor& cellProcessor) override;
	void traversePartialInnermostCells(CellProcessor& cellProcessor, unsigned int stage, int stageCount) override;
	void deleteOuterParticles() override;                                   // LinkedCells.cpp:611-628
	unsigned long getNumberOfParticles() override;                          // LinkedCells.cpp:577-588
	void clear() override;                                                  // LinkedCells.cpp:590-595
	bool requiresForceExchange() const override { return false; }           // full shell: halo forces are never sent back
	std::string getName() override { return "GpuLinkedCells"; }
	ParticleIterator iterator(ParticleIterator::Type t) override;          // LinkedCells.h:245-250
	RegionParticleIterator regionIterator(const double startRegion[3], const double endRegion[3],
	                                      ParticleIterator::Type type) override;

	// called by GpuLeapfrog (Integrator::eventNewTimestep / eventForcesCalculated)
	void integratePre(double dt, Domain* domain);                           // Leapfrog.cpp:48-68
	void integratePost(double dt, Domain* domain);                          // Leapfrog.cpp:70-155

	/** refill the host cells from the device (r, v, q, D, F, M, Vi); they are dropped again at the next time step */
	void downloadMirror();
	/** true while the device, not the host cells, holds the current state (device phase, outside a time step, no mirror taken) */
	bool deviceHoldsState() const { return _ctx != nullptr && _deviceLoaded && _devicePhase && !_mirrorValid && !_inStep; }
	/** the owned molecules of all blocks straight from the device into `file`, block after block as
	 *  DomainDecompBase::writeMoleculesToFile appends rank after rank, with the velocity scaling of the step that just ended applied
	 *  on the way out as VelocityScalingThermostat::apply would have by now.
	 *  ascii = false: FullMolecule::writeBinary records (ICRVQD, 116 bytes); the file is truncated first.
	 *  ascii = true : FullMolecule::write lines at the writer's stream precision of 20 (RecordText.h), formatted by the host threads
	 *                 from the packed records; appended to the file, which holds Domain::writeCheckpointHeader's header already.
	 *  Requires deviceHoldsState().  Returns the number of molecules written. */
	uint64_t writeRecords(const std::string& file, Domain* domain, bool ascii);
	ls1g_ctx* ctx() { return _ctx; }
	int numDevices() const { return static_cast<int>(_ctxs.size()); }

private:
	enum Mode { FUSED, OPERATORS };
	void ensureContext();
	void uploadComponents();
	void uploadFromHost();
	void uploadBlocks(const std::vector<size_t>& start);
	int blockOfCell(const ParticleCell& pc) const;
	void downloadMirrorBlocks();
	int blockOf(const double* r) const;
	double ensureStaging(size_t molecules);
	// one call for the whole box, whatever the number of blocks
	void exchangeAll();
	void traverseAll(double& upot, double& virial);
	bool multi() const { return _ctxs.size() > 1; }
	void clearHostCells();
	bool mirrorWanted() const;
	unsigned long mirrorFrequencyFromPlugins(XMLfileUnits& xmlconfig) const;
	void printSummary() const;
	void applySelectiveThermostat(Domain* domain);
	void sampleRDF();
	void sampleFlops(FlopCounter& counter);
	[[noreturn]] void fail(const char* what) const;
	void chk(int rc, const char* what) const;

	ls1g_ctx* _ctx = nullptr;                 // block 0 (the only one unless <devices> > 1)
	std::vector<ls1g_ctx*> _ctxs;            // all blocks in rank order
	int _blockGrid[3] = {1, 1, 1};
	std::vector<double> _blockLo[3];         // lower bound of block k in dimension d (ls1g_decomp_box), + the upper end
	Mode _mode = FUSED;
	int _device = 0, _devices = 1;
	bool _allowUnknownPlugins = false;
	unsigned long _mirrorFrequency = 0;
	bool _devicePhase = false;     // false: the host cells hold the molecules (before the first time step)
	bool _deviceLoaded = false;    // the device holds a copy with valid forces
	bool _mirrorValid = false;     // host cells hold a snapshot of the device state
	bool _inStep = false;          // between integratePre and integratePost
	bool _ranThisStep = false;     // fused mode: the step already ran inside traverseCells
	double _dt = 0.;
	double _summv2 = 0., _sumIw2 = 0.;
	uint64_t _n = 0, _rotDOF = 0;
	// wall clock of the host<->device phases (logged after the last time step)
	double _tUpload = 0., _tDevice = 0., _tMirror = 0., _tContext = 0.;
	double _tGather = 0., _tH2D = 0., _tD2H = 0., _tFill = 0.;
	bool _pointMolecules = false;  // one component, one LJ centre at the origin, no rotation: the device keeps no q / D / M
	unsigned long _steps = 0, _mirrors = 0, _selective = 0, _rdfSamples = 0, _flopSamples = 0;
	double _lastTfactor = 1.0;
	unsigned long _selectiveStep = ~0ul;   // value of _steps for which the explosion heuristics have been looked at (once per step)
	unsigned long _recordWrites = 0;
	double _tRecords = 0.;
	size_t _bytesRecords = 0;
	size_t _bytesUp = 0, _bytesDown = 0, _numCells = 0;
	double _pendingBt = 1.0, _pendingBr = 1.0;   // fused mode: betas of the previous step, applied inside the next ls1g_step
	// page-locked staging block (ls1g_host_alloc): r v q L F M Vi id cid cell, 192 bytes per molecule
	static constexpr size_t kStageBytesPerMolecule = 24 * 6 + 32 + 8 + 4 + 4;
	void* _stage = nullptr;
	bool _stagePinned = true;
	double _tFirstSteps[4] = {0., 0., 0., 0.}, _tMaxStep = 0.;
	size_t _stageMolecules = 0;
	double *_sR = nullptr, *_sV = nullptr, *_sQ = nullptr, *_sL = nullptr, *_sF = nullptr, *_sM = nullptr, *_sVi = nullptr;
	uint64_t* _sId = nullptr;
	uint32_t *_sCid = nullptr, *_sCell = nullptr;
};

#endif /* GPULINKEDCELLS_H_ */
