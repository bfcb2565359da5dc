// Shared between ensemble_io.cu (record semantics of strucreader.f90) and ensemble_fast.cu (line index +
// device conversion of the atom records).
#pragma once
#include <cstddef>
#include <vector>

namespace crest {

struct MappedFile {
  int fd = -1;
  const char* data = nullptr;
  size_t size = 0;
  ~MappedFile();
  int open_read(const char* fname);     // 0 ok (an empty file maps to data = NULL, size = 0)
};

void index_lines(const char* data, size_t size, std::vector<unsigned long long>& line_start);

// Converts the atom records of every structure (first_line[s] .. + nat) on the current device into
// xyz_host[s][j][0..2]; flags[s*nat + j]: 0 converted, 1 the host must re-read the record (a number beyond
// the exact fast path), 2 the record does not scan.  validate_only: no output, flag 2 unless the record
// reads as `str x y z` (rdensembleparam).  Returns 0 ok, 1 CUDA error (text set).
int parse_atoms_device(const MappedFile& f, const std::vector<unsigned long long>& line_start,
                       const std::vector<long long>& first_line, int nat, double* xyz_host,
                       std::vector<unsigned>& flags, int validate_only);

}  // namespace crest
