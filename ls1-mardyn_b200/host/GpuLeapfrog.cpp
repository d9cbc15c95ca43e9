#include "GpuLeapfrog.h"

#include "GpuLinkedCells.h"
#include "Simulation.h"
#include "utils/Logger.h"
#include "utils/xmlfileUnits.h"

using Log::global_log;

namespace {
GpuLinkedCells* deviceContainer(ParticleContainer* pc) {
	GpuLinkedCells* g = dynamic_cast<GpuLinkedCells*>(pc);
	if (g == nullptr) {
		global_log->error_always_output() << "GpuLeapfrog needs <datastructure type=\"GpuLinkedCells\">" << std::endl;
		Simulation::exit(1);
	}
	return g;
}
}  // namespace

void GpuLeapfrog::readXML(XMLfileUnits& xmlconfig) {
	_timestepLength = 0;
	xmlconfig.getNodeValueReduced("timestep", _timestepLength);
	global_log->info() << "Timestep: " << _timestepLength << std::endl;
}

void GpuLeapfrog::eventNewTimestep(ParticleContainer* moleculeContainer, Domain* domain) {
	deviceContainer(moleculeContainer)->integratePre(_timestepLength, domain);
}

void GpuLeapfrog::eventForcesCalculated(ParticleContainer* moleculeContainer, Domain* domain) {
	deviceContainer(moleculeContainer)->integratePost(_timestepLength, domain);
}
