/*
 * GpuCheckpointWriter — ls1-MarDyn output plugin: binary checkpoints written straight from the device.
 *
 *   <outputplugin name="GpuCheckpointWriter">
 *     <type>binary|ASCII</type>                    default binary
 *     <writefrequency>INTEGER</writefrequency>     default 1
 *     <outputprefix>STRING</outputprefix>          default "mardyn"
 *     <incremental>INTEGER</incremental>           default 1: PREFIX-<number>.restart.{header.xml,dat} (binary), PREFIX-<number>.restart.dat (ASCII)
 *   </outputplugin>
 *
 * Same files as the reference's <outputplugin name="CheckpointWriter"><type>binary</type> (io/CheckpointWriter.cpp:62-90 ->
 * Domain::writeCheckpoint, Domain.cpp:597-619): the header comes from the reference's own Domain::writeCheckpointHeaderXML, the
 * .dat file holds one FullMolecule::writeBinary record (molecules/FullMolecule.cpp:451-474) per molecule.  The difference is where
 * the records come from: DomainDecompBase::writeMoleculesToFile (parallel/DomainDecompBase.cpp:505-540, not virtual) walks host
 * Molecule objects, which costs a GpuLinkedCells container a full mirror step; this plugin asks the container for the packed records
 * (GpuLinkedCells::writeRecords -> ls1g_download_records) and writes the bytes as they arrive - no host Molecule is touched.
 * <type>ASCII</type> gives the files of <CheckpointWriter><type>ASCII</type>: the reference's Domain::writeCheckpointHeader, then one
 * FullMolecule::write line per molecule, printed by the host threads from the packed records (RecordText.h).
 * The molecules come cell by cell as in the reference's file, inside a cell ordered by id.  With any other container, or while
 * the host cells hold the state, it falls back to Domain::writeCheckpoint(binary).
 *
 * Registered by one line next to the default plugins:
 *     pluginFactory.registerPlugin(&GpuCheckpointWriter::createInstance);
 */
#ifndef GPUCHECKPOINTWRITER_H_
#define GPUCHECKPOINTWRITER_H_

#include <string>

#include "plugins/PluginBase.h"

class GpuCheckpointWriter : public PluginBase {
public:
	void readXML(XMLfileUnits& xmlconfig) override;
	void init(ParticleContainer*, DomainDecompBase*, Domain*) override {}
	void endStep(ParticleContainer* particleContainer, DomainDecompBase* domainDecomp, Domain* domain, unsigned long simstep) override;
	void finish(ParticleContainer*, DomainDecompBase*, Domain*) override {}
	std::string getPluginName() override { return "GpuCheckpointWriter"; }
	static PluginBase* createInstance() { return new GpuCheckpointWriter(); }

private:
	std::string _prefix = "mardyn";
	unsigned long _every = 1;
	bool _numbered = true;
	bool _ascii = false;
};

#endif /* GPUCHECKPOINTWRITER_H_ */
