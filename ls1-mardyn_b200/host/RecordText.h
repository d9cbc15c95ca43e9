/*
 * RecordText.h — the ASCII form of a phase-space record, produced from the packed binary record without a Molecule object.
 *
 * One line per molecule as FullMolecule::write prints it (molecules/FullMolecule.cpp:442-449) into the stream that
 * DomainDecompBase::writeMoleculesToFile opened with precision(20) (parallel/DomainDecompBase.cpp:521-523):
 *     id \t componentid \t x y z \t vx vy vz \t qw qx qy qz \t Dx Dy Dz \t \n
 * An ostream in its default float format with precision 20 prints a double as printf's %.20g does (libstdc++'s num_put formats
 * through vsnprintf with exactly that conversion), so the lines are byte for byte the reference's.  tests/test_records.py holds
 * this file to a checkpoint pair the reference wrote (binary and ASCII of the same state).
 */
#ifndef RECORDTEXT_H_
#define RECORDTEXT_H_

#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

namespace ls1g_records {

constexpr size_t kRecordBytes = 116;      // FullMolecule::writeBinary: 8 + 4 + 13 * 8
constexpr size_t kMaxLineBytes = 512;     // 2 integers + 13 doubles of at most 28 characters + separators

/** appends the line of one ICRVQD record (packed, unaligned) to `out`; returns the number of characters written */
inline size_t appendLine(const unsigned char* rec, char* out) {
	uint64_t id;
	uint32_t cid;
	double d[13];
	std::memcpy(&id, rec, 8);
	std::memcpy(&cid, rec + 8, 4);
	std::memcpy(d, rec + 12, 13 * 8);
	const int n = std::snprintf(out, kMaxLineBytes,
	                            "%llu\t%u\t%.20g %.20g %.20g\t%.20g %.20g %.20g\t%.20g %.20g %.20g %.20g\t%.20g %.20g %.20g\t\n",
	                            static_cast<unsigned long long>(id), cid, d[0], d[1], d[2], d[3], d[4], d[5], d[6], d[7], d[8], d[9],
	                            d[10], d[11], d[12]);
	return n > 0 ? static_cast<size_t>(n) : 0;
}

/** the lines of records [first, last) of a packed buffer */
inline void appendLines(const unsigned char* records, size_t first, size_t last, std::string& text) {
	char line[kMaxLineBytes];
	for (size_t i = first; i < last; ++i) text.append(line, appendLine(records + i * kRecordBytes, line));
}

}  // namespace ls1g_records

#endif /* RECORDTEXT_H_ */
