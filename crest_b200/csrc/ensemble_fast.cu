// ensemble_fast.cu — the xyz-ensemble reader at BASELINE sizes.
//
// A C3 ensemble is a 0.4 GB text file (C5: 2.7 GB) and the CREGEN call behind it takes ~20 ms; the
// record-by-record reader of ensemble_io.cu (the restatement of strucreader.f90:268-309, 433-481) needs
// ~10 s for it, like the reference's list-directed READs.  Same results, three stages:
//   1. host: mmap the file, index its lines with a few threads (memchr);
//   2. host: walk the structures exactly like the record reader (count record with blank records
//      skipped and non-integer counts ignored, comment record -> grepenergy, `nat` atom records,
//      ensemble_strucskip for foreign atom counts) - integer logic on ~2 records per structure;
//   3. device: ONE THREAD PER ATOM RECORD converts the three coordinates (coordline, strucreader.f90:
//      2005-2022: "x y z sym" first, then "sym x y z") with the exact fast path of decimal -> binary64
//      conversion: <= 19 significant digits collected in an integer that fits 2^53 and a decimal
//      exponent within +-22 make  double(m) (* | /) 10^k  ONE correctly rounded operation, i.e. bit-identical
//      to strtod / a correctly rounding READ.  Anything else (more digits, huge exponents, a record
//      that does not scan) is flagged and re-read by the record reader's own code on the host.
// The text crosses PCIe once (staged from the page cache), the doubles come back once.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <algorithm>
#include <cstring>
#include <string>
#include <thread>
#include <vector>
#include "common.cuh"
#include "hostcopy.cuh"
#include "ensemble_fast.cuh"

namespace crest {

namespace {

__device__ __constant__ double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                             1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

__device__ __forceinline__ bool is_sep(char c) { return c == ' ' || c == '\t' || c == '\r'; }
__device__ __forceinline__ bool is_dig(char c) { return c >= '0' && c <= '9'; }

// next list-directed item of [p, e): returns false at the end of the record (ensemble_io.cu next_item)
__device__ bool next_item_dev(const char*& p, const char* e, const char*& b, const char*& t) {
  while (p < e && is_sep(*p)) ++p;
  if (p >= e || *p == '/') return false;
  b = p;
  while (p < e && !is_sep(*p) && *p != ',' && *p != '/') ++p;
  t = p;
  const char* q = p;
  while (q < e && (*q == ' ' || *q == '\t')) ++q;
  if (q < e && *q == ',') p = q + 1;
  return true;
}

// Fortran real item (ensemble_io.cu parse_real).  0 = not a real, 1 = converted exactly, 2 = a real the
// fast path cannot convert exactly (host re-reads the record)
__device__ int parse_real_dev(const char* b, const char* e, double* out) {
  const char* p = b;
  bool neg = false;
  if (p < e && (*p == '+' || *p == '-')) { neg = *p == '-'; ++p; }
  unsigned long long m = 0;
  int nd = 0, drop = 0, frac = 0;
  bool digits = false, slow = false;
  while (p < e && is_dig(*p)) {
    digits = true;
    if (m || *p != '0') { if (nd < 19) { m = m * 10 + (unsigned)(*p - '0'); ++nd; } else { ++drop; slow = true; } }
    ++p;
  }
  if (p < e && *p == '.') {
    ++p;
    while (p < e && is_dig(*p)) {
      digits = true;
      if (m || *p != '0') { if (nd < 19) { m = m * 10 + (unsigned)(*p - '0'); ++nd; ++frac; } else slow = true; }
      else ++frac;
      ++p;
    }
  }
  if (!digits) return 0;
  int ex = 0;
  if (p < e) {
    bool eneg = false;
    if (*p == 'e' || *p == 'E' || *p == 'd' || *p == 'D') {
      ++p;
      if (p < e && (*p == '+' || *p == '-')) { eneg = *p == '-'; ++p; }
    } else if (*p == '+' || *p == '-') {
      eneg = *p == '-'; ++p;
    } else {
      return 0;
    }
    bool ed = false;
    while (p < e && is_dig(*p)) { ed = true; if (ex < 100000) ex = ex * 10 + (*p - '0'); ++p; }
    if (!ed || p != e) return 0;
    if (eneg) ex = -ex;
  }
  const int k = ex - frac + drop;
  if (slow || m > (1ull << 53) || k < -22 || k > 22) return 2;
  double v = (double)(long long)m;                    // exact: m <= 2^53
  v = k >= 0 ? v * kPow10[k] : v / kPow10[-k];        // one correctly rounded operation
  *out = neg ? -v : v;
  return 1;
}

struct ParseArgs {
  const char* text;
  const unsigned long long* line_start;     // [nlines + 1]
  const long long* first_line;              // per structure: index of its first atom record
  int nat, nstruct;
  double* xyz;                              // [nstruct][nat][3]
  unsigned* flags;                          // per atom record: 0 ok, 1 host must re-read, 2 does not scan
  int validate_only;                        // rdensembleparam: the record must read as `str x y z`
};

__global__ void __launch_bounds__(256)
parse_atoms_kernel(const ParseArgs a) {
  const long long tid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (tid >= (long long)a.nstruct * a.nat) return;
  const int s = (int)(tid / a.nat), j = (int)(tid % a.nat);
  const long long ln = a.first_line[s] + j;
  const char* p = a.text + a.line_start[ln];
  const char* e = a.text + a.line_start[ln + 1];
  if (e > p && e[-1] == '\n') --e;
  if (e > p && e[-1] == '\r') --e;
  const char *b[4], *t[4];
  int n = 0;
  while (n < 4 && next_item_dev(p, e, b[n], t[n])) ++n;
  unsigned flag = 0;
  double v[3] = {0.0, 0.0, 0.0};
  if (n < 4) {
    flag = 2;
  } else if (a.validate_only) {
    for (int k = 0; k < 3; ++k) if (parse_real_dev(b[k + 1], t[k + 1], &v[k]) == 0) flag = 2;
  } else {
    int r0 = parse_real_dev(b[0], t[0], &v[0]);
    int r1 = r0 ? parse_real_dev(b[1], t[1], &v[1]) : 0;
    int r2 = r1 ? parse_real_dev(b[2], t[2], &v[2]) : 0;
    if (!(r0 && r1 && r2)) {
      r0 = parse_real_dev(b[1], t[1], &v[0]);
      r1 = r0 ? parse_real_dev(b[2], t[2], &v[1]) : 0;
      r2 = r1 ? parse_real_dev(b[3], t[3], &v[2]) : 0;
    }
    if (!(r0 && r1 && r2)) flag = 2;
    else if (r0 == 2 || r1 == 2 || r2 == 2) flag = 1;
  }
  a.flags[tid] = flag;
  if (!a.validate_only) {
    double* dst = a.xyz + tid * 3;
    dst[0] = v[0]; dst[1] = v[1]; dst[2] = v[2];
  }
}

struct DevBufF {
  void* p = nullptr;
  ~DevBufF() { cudaFree(p); }
  int get(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1) == cudaSuccess ? 0 : 1; }
};

}  // namespace

MappedFile::~MappedFile() {
  if (data && size) munmap(const_cast<char*>(data), size);
  if (fd >= 0) close(fd);
}

int MappedFile::open_read(const char* fname) {
  fd = ::open(fname, O_RDONLY);
  if (fd < 0) return 1;
  struct stat st;
  if (fstat(fd, &st) != 0) return 1;
  size = (size_t)st.st_size;
  if (size == 0) { data = nullptr; return 0; }
  void* m = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
  if (m == MAP_FAILED) { size = 0; return 1; }
  madvise(m, size, MADV_SEQUENTIAL);
  data = (const char*)m;
  return 0;
}

// line_start[k] = offset of record k, line_start[nlines] = size (a last record without '\n' counts)
void index_lines(const char* data, size_t size, std::vector<unsigned long long>& line_start) {
  line_start.clear();
  if (size == 0) { line_start.push_back(0); return; }
  int T = (int)std::min<size_t>(8, std::max<size_t>(1, size >> 24));
  const int hw = (int)std::thread::hardware_concurrency();
  if (hw > 0 && T > hw) T = hw;
  std::vector<std::vector<unsigned long long>> part((size_t)T);
  auto scan = [&](int t) {
    const size_t b = size * t / T, e = size * (t + 1) / T;
    std::vector<unsigned long long>& v = part[t];
    v.reserve((e - b) / 40 + 16);
    const char* p = data + b;
    const char* end = data + e;
    while (p < end) {
      const char* q = (const char*)memchr(p, '\n', (size_t)(end - p));
      if (!q) break;
      v.push_back((unsigned long long)(q - data) + 1);   // the NEXT record starts behind the newline
      p = q + 1;
    }
  };
  std::vector<std::thread> th;
  for (int t = 1; t < T; ++t) th.emplace_back(scan, t);
  scan(0);
  for (auto& x : th) x.join();
  size_t n = 1;
  for (auto& v : part) n += v.size();
  line_start.reserve(n + 1);
  line_start.push_back(0);
  for (auto& v : part) line_start.insert(line_start.end(), v.begin(), v.end());
  if (line_start.back() != size) line_start.push_back(size);
}

int parse_atoms_device(const MappedFile& f, const std::vector<unsigned long long>& line_start,
                       const std::vector<long long>& first_line, int nat, double* xyz_host,
                       std::vector<unsigned>& flags, int validate_only) {
  const size_t ns = first_line.size();
  flags.assign(ns * (size_t)nat, 0u);
  if (ns == 0) return 0;
  int dev = 0;
  CREST_CUDA_TRY(cudaGetDevice(&dev));
  DevBufF d_text, d_ls, d_fl, d_xyz, d_flags;
  if (d_text.get(f.size) || d_ls.get(sizeof(unsigned long long) * line_start.size()) ||
      d_fl.get(sizeof(long long) * ns) || d_flags.get(sizeof(unsigned) * ns * nat) ||
      (!validate_only && d_xyz.get(sizeof(double) * 3 * nat * ns))) {
    set_error("cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
    return 1;
  }
  if (big_copy(d_text.p, f.data, f.size, 0, dev)) return 1;
  if (big_copy(d_ls.p, line_start.data(), sizeof(unsigned long long) * line_start.size(), 0, dev)) return 1;
  CREST_CUDA_TRY(cudaMemcpy(d_fl.p, first_line.data(), sizeof(long long) * ns, cudaMemcpyHostToDevice));
  ParseArgs a{};
  a.text = (const char*)d_text.p; a.line_start = (const unsigned long long*)d_ls.p;
  a.first_line = (const long long*)d_fl.p; a.nat = nat; a.nstruct = (int)ns;
  a.xyz = (double*)d_xyz.p; a.flags = (unsigned*)d_flags.p; a.validate_only = validate_only;
  const long long total = (long long)ns * nat;
  parse_atoms_kernel<<<(unsigned)((total + 255) / 256), 256>>>(a);
  CREST_CUDA_TRY(cudaGetLastError());
  CREST_CUDA_TRY(cudaDeviceSynchronize());
  CREST_CUDA_TRY(cudaMemcpy(flags.data(), d_flags.p, sizeof(unsigned) * flags.size(), cudaMemcpyDeviceToHost));
  if (!validate_only && big_copy(xyz_host, d_xyz.p, sizeof(double) * 3 * nat * ns, 1, dev)) return 1;
  return 0;
}

}  // namespace crest
