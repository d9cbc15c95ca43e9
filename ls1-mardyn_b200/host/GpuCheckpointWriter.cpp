/*
 * GpuCheckpointWriter.cpp — see GpuCheckpointWriter.h.
 */
#include "GpuCheckpointWriter.h"

#include "GpuLinkedCells.h"

#include "Common.h"
#include "Domain.h"
#include "Simulation.h"
#include "parallel/DomainDecompBase.h"
#include "utils/Logger.h"
#include "utils/xmlfileUnits.h"

#include <cmath>

using Log::global_log;

void GpuCheckpointWriter::readXML(XMLfileUnits& xmlconfig) {
	long every = 1;
	xmlconfig.getNodeValue("writefrequency", every);
	if (every <= 0) {
		global_log->error() << "GpuCheckpointWriter: writefrequency must be a positive integer, but is " << every << std::endl;
		Simulation::exit(-1);
	}
	_every = static_cast<unsigned long>(every);
	std::string type = "binary";
	xmlconfig.getNodeValue("type", type);
	if (type != "binary" && type != "ASCII") {
		global_log->error() << "GpuCheckpointWriter: unknown type '" << type << "', expected: ASCII|binary." << std::endl;
		Simulation::exit(-1);
	}
	_ascii = type == "ASCII";
	xmlconfig.getNodeValue("outputprefix", _prefix);
	int numbered = 1;
	xmlconfig.getNodeValue("incremental", numbered);
	_numbered = numbered != 0;
	global_log->info() << "GpuCheckpointWriter: " << type << " checkpoints every " << _every << " step(s), prefix " << _prefix
	                   << (_numbered ? ", numbered" : "") << std::endl;
}

void GpuCheckpointWriter::endStep(ParticleContainer* particleContainer, DomainDecompBase* domainDecomp, Domain* domain,
                                  unsigned long simstep) {
	if (simstep % _every != 0) return;
	// the reference's file names (io/CheckpointWriter.cpp:66-84): PREFIX[-<step / frequency>, zero-padded].restart (binary: + .dat / .header.xml; ASCII: .restart.dat)
	std::string name = _prefix;
	if (_numbered) {
		const unsigned long writes = _simulation.getNumTimesteps() / _every;
		const int digits = writes >= 1 ? static_cast<int>(std::ceil(std::log(static_cast<double>(writes)) / std::log(10.))) : 0;
		name += "-" + aligned_number(static_cast<int>(simstep / _every), digits, '0');
	}
	name += _ascii ? ".restart.dat" : ".restart";
	const double time = _simulation.getSimulationTime();
	auto* gpu = dynamic_cast<GpuLinkedCells*>(particleContainer);
	if (gpu == nullptr || !gpu->deviceHoldsState()) {
		domain->writeCheckpoint(name, particleContainer, domainDecomp, time, !_ascii);
		return;
	}
	domain->updateglobalNumMolecules(particleContainer, domainDecomp);                                  // Domain.cpp:611
	if (_ascii) {
		domain->writeCheckpointHeader(name, particleContainer, domainDecomp, time);                     // Domain.cpp:616: header into the same file
		gpu->writeRecords(name, domain, true);
	} else {
		domain->writeCheckpointHeaderXML(name + ".header.xml", particleContainer, domainDecomp, time);  // Domain.cpp:614
		gpu->writeRecords(name + ".dat", domain, false);
	}
}
