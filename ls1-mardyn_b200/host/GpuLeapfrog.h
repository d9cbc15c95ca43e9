/*
 * GpuLeapfrog — Integrator (src/integrators/Integrator.h:33-90) that forwards the two Leapfrog events
 * (src/integrators/Leapfrog.cpp:42-155) to the device context of a GpuLinkedCells container.
 * Selected with <integrator type="GpuLeapfrog"><timestep unit="reduced">...</timestep></integrator>.
 */
#ifndef GPULEAPFROG_H_
#define GPULEAPFROG_H_

#include "integrators/Integrator.h"

class GpuLeapfrog : public Integrator {
public:
	GpuLeapfrog() : Integrator() {}
	explicit GpuLeapfrog(double timestepLength) : Integrator(timestepLength) {}
	~GpuLeapfrog() override {}

	void readXML(XMLfileUnits& xmlconfig) override;                                              // Leapfrog.cpp:24-29
	void init() override {}
	void eventNewTimestep(ParticleContainer* moleculeContainer, Domain* domain) override;        // Leapfrog.cpp:48-68
	void eventForcesCalculated(ParticleContainer* moleculeContainer, Domain* domain) override;   // Leapfrog.cpp:70-155
};

#endif /* GPULEAPFROG_H_ */
