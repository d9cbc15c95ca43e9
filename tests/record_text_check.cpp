// Test driver for ls1-mardyn_b200/host/RecordText.h (tests/test_records.py compiles it): packed ICRVQD records of a .dat file
// in, the ASCII molecule lines the product writes for them out.
#include "RecordText.h"

#include <fstream>
#include <iostream>
#include <iterator>
#include <vector>

int main(int argc, char** argv) {
	if (argc != 2) return 2;
	std::ifstream in(argv[1], std::ios::binary);
	std::vector<unsigned char> bytes((std::istreambuf_iterator<char>(in)), std::istreambuf_iterator<char>());
	if (bytes.size() % ls1g_records::kRecordBytes != 0) return 3;
	std::string text;
	ls1g_records::appendLines(bytes.data(), 0, bytes.size() / ls1g_records::kRecordBytes, text);
	std::cout << text;
	return 0;
}
