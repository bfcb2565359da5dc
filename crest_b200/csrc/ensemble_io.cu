// ensemble_io.cu — host-only: the xyz-ensemble reader / writer either side of CREGEN.
//
//   rdensembleparam   /root/reference/src/strucreader.f90:268-309   (count structures, largest nat)
//   rdensemble_conf3  /root/reference/src/strucreader.f90:433-481   (at, xyz, comment lines)
//   coordline         /root/reference/src/strucreader.f90:2005-2022 ("x y z sym" first, then "sym x y z")
//   e2i / i2e         /root/reference/src/strucreader.f90:2066-2176 (label -> Z, Z -> "nice case" symbol)
//   grepenergy        /root/reference/src/strucreader.f90:2225-2260 (energy=/energy: keyword, else first float)
//   wrxyz_channel     /root/reference/src/strucreader.f90:1679-1698 ('(2x,i0)', comment, '(1x,a2,1x,3f20.10)')
//
// The reference reads with Fortran list-directed input.  What is reproduced: blank records before
// an item are skipped, items are separated by blanks / commas, reals accept d/D/e/E exponents and
// the exponent-letter-free form (1.5+3), an integer item must be an integer, a failed count line is
// skipped (`cycle`), a structure with a different atom count is skipped by the reader
// (ensemble_strucskip) and a structure with an unreadable atom line is not counted.  Not reproduced:
// an atom record continued over several lines, repeat counts (3*1.0) and quoted strings — no xyz
// writer produces them.  Text parsing stays on the host: a 100 000 x 60 ensemble is 0.4 GB of text,
// read once per CREGEN call, and the parsed doubles go straight into the array the C ABI uploads.
#include <cctype>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "common.cuh"
#include "ensemble_fast.cuh"
#include "../../include/crest_cuda.h"

namespace {

using crest::set_error;

const char* const kPSE[118] = {
    "H",  "He", "Li", "Be", "B",  "C",  "N",  "O",  "F",  "Ne", "Na", "Mg", "Al", "Si", "P",  "S",
    "Cl", "Ar", "K",  "Ca", "Sc", "Ti", "V",  "Cr", "Mn", "Fe", "Co", "Ni", "Cu", "Zn", "Ga", "Ge",
    "As", "Se", "Br", "Kr", "Rb", "Sr", "Y",  "Zr", "Nb", "Mo", "Tc", "Ru", "Rh", "Pd", "Ag", "Cd",
    "In", "Sn", "Sb", "Te", "I",  "Xe", "Cs", "Ba", "La", "Ce", "Pr", "Nd", "Pm", "Sm", "Eu", "Gd",
    "Tb", "Dy", "Ho", "Er", "Tm", "Yb", "Lu", "Hf", "Ta", "W",  "Re", "Os", "Ir", "Pt", "Au", "Hg",
    "Tl", "Pb", "Bi", "Po", "At", "Rn", "Fr", "Ra", "Ac", "Th", "Pa", "U",  "Np", "Pu", "Am", "Cm",
    "Bk", "Cf", "Es", "Fm", "Md", "No", "Lr", "Rf", "Db", "Sg", "Bh", "Hs", "Mt", "Ds", "Rg", "Cn",
    "Nh", "Fl", "Mc", "Lv", "Ts", "Og"};

// one list-directed item: [begin, end) of the next token of `s` starting at pos, or false at the end
// of the record.  Separators: blanks, tabs, one comma; '/' ends the record.
bool next_item(const std::string& s, size_t& pos, size_t& b, size_t& e) {
  while (pos < s.size() && (s[pos] == ' ' || s[pos] == '\t' || s[pos] == '\r')) ++pos;
  if (pos >= s.size() || s[pos] == '/') return false;
  b = pos;
  while (pos < s.size() && s[pos] != ' ' && s[pos] != '\t' && s[pos] != ',' && s[pos] != '\r' &&
         s[pos] != '/') ++pos;
  e = pos;
  size_t q = pos;
  while (q < s.size() && (s[q] == ' ' || s[q] == '\t')) ++q;
  if (q < s.size() && s[q] == ',') pos = q + 1;
  return true;
}

// Fortran real item: [sign] digits [. digits] [ (e|E|d|D) [sign] digits | sign digits ]
bool parse_real(const char* b, const char* e, double* out) {
  std::string t;
  const char* p = b;
  if (p < e && (*p == '+' || *p == '-')) t += *p++;
  bool digits = false;
  while (p < e && isdigit((unsigned char)*p)) { t += *p++; digits = true; }
  if (p < e && *p == '.') {
    t += *p++;
    while (p < e && isdigit((unsigned char)*p)) { t += *p++; digits = true; }
  }
  if (!digits) return false;
  if (p < e) {
    if (*p == 'e' || *p == 'E' || *p == 'd' || *p == 'D') {
      ++p;
      t += 'e';
      if (p < e && (*p == '+' || *p == '-')) t += *p++;
    } else if (*p == '+' || *p == '-') {
      t += 'e';
      t += *p++;
    } else {
      return false;
    }
    bool ed = false;
    while (p < e && isdigit((unsigned char)*p)) { t += *p++; ed = true; }
    if (!ed || p != e) return false;
  }
  *out = strtod(t.c_str(), nullptr);
  return true;
}

bool parse_int(const char* b, const char* e, long* out) {
  const char* p = b;
  if (p < e && (*p == '+' || *p == '-')) ++p;
  if (p == e) return false;
  for (const char* q = p; q < e; ++q) if (!isdigit((unsigned char)*q)) return false;
  *out = strtol(std::string(b, e).c_str(), nullptr, 10);
  return true;
}

// convertlable (strucreader.f90:2066-2090): isotope / charge decorations become blanks, upper case,
// left adjusted, second letter lower case
std::string convert_label(const std::string& s) {
  std::string o = s;
  for (char& c : o) {
    if (strchr("0123456789*_+-", c)) c = ' ';
    else c = (char)toupper((unsigned char)c);
  }
  size_t b = o.find_first_not_of(' ');
  if (b == std::string::npos) return "";
  size_t e = o.find_last_not_of(' ');
  o = o.substr(b, e - b + 1);
  if (o.size() > 1) o[1] = (char)tolower((unsigned char)o[1]);
  return o;
}

int e2i(const std::string& sym) {   // strucreader.f90:2120-2147
  const std::string c = convert_label(sym);
  for (int i = 0; i < 118; ++i) if (c == kPSE[i]) return i + 1;
  long z;
  if (parse_int(sym.data(), sym.data() + sym.size(), &z) && z <= 118) return (int)z;
  if (c == "D" || c == "T") return 1;
  return 0;
}

// coordline (strucreader.f90:2005-2022): "x y z sym", else "sym x y z"
bool coordline(const std::string& line, std::string* sym, double* xyz) {
  size_t b[4], e[4], pos = 0;
  int n = 0;
  while (n < 4 && next_item(line, pos, b[n], e[n])) ++n;
  if (n < 4) return false;
  const char* s = line.data();
  if (parse_real(s + b[0], s + e[0], xyz) && parse_real(s + b[1], s + e[1], xyz + 1) &&
      parse_real(s + b[2], s + e[2], xyz + 2)) {
    *sym = line.substr(b[3], e[3] - b[3]);
    return true;
  }
  if (parse_real(s + b[1], s + e[1], xyz) && parse_real(s + b[2], s + e[2], xyz + 1) &&
      parse_real(s + b[3], s + e[3], xyz + 2)) {
    *sym = line.substr(b[0], e[0] - b[0]);
    return true;
  }
  return false;
}

// Records of a memory-mapped file through its line index (ensemble_fast.cu): what `read(ich,'(a)')`
// delivers, one record per call, '\r' of DOS files dropped.
struct LineReader {
  const char* data;
  const std::vector<unsigned long long>& ls;
  size_t k = 0;
  LineReader(const crest::MappedFile& f, const std::vector<unsigned long long>& index) : data(f.data), ls(index) {}
  size_t nlines() const { return ls.size() - 1; }
  bool next(std::string& out) {
    if (k >= nlines()) return false;
    size_t b = ls[k], e = ls[k + 1];
    ++k;
    if (e > b && data[e - 1] == '\n') --e;
    if (e > b && data[e - 1] == '\r') --e;
    out.assign(data + b, e - b);
    return true;
  }
  bool skip(size_t n) {                      // n records without looking at them; false at end of file
    if (k + n > nlines()) { k = nlines(); return false; }
    k += n;
    return true;
  }
};

// atom records are converted on the device when there is one and the file is large enough to pay for the
// transfer of its text (ensemble_fast.cu); the host loop below is the same coordline() either way
bool use_device_parser(size_t natom_records) {
  if (natom_records < 200000) return false;
  const char* e = getenv("CREST_CUDA_IO");
  if (e && !strcmp(e, "host")) return false;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return false; }
  return n > 0;
}

// `read(ich,*,iostat=io) dum`: skip empty records, then the first item must be an integer.
// returns -1 end of file, 1 conversion error (record consumed), 0 ok
int read_count(LineReader& R, long* dum) {
  std::string line;
  for (;;) {
    if (!R.next(line)) return -1;
    size_t pos = 0, b, e;
    if (!next_item(line, pos, b, e)) continue;         // blank record: keep looking
    return parse_int(line.data() + b, line.data() + e, dum) ? 0 : 1;
  }
}

double grepenergy(const std::string& raw) {             // strucreader.f90:2225-2260
  std::string a = raw;
  while (!a.empty() && (a.back() == ' ' || a.back() == '\r' || a.back() == '\n')) a.pop_back();
  auto read_first = [](const std::string& s, double* v) -> int {   // read(atmp,*,iostat=io) energy
    size_t pos = 0, b, e;
    size_t q = 0;
    while (q < s.size() && (s[q] == ' ' || s[q] == '\t')) ++q;
    if (q >= s.size()) return -1;                       // nothing to read: end of record
    if (s[q] == ',' || s[q] == '/') return 0;           // null value: energy keeps its 0.0
    if (!next_item(s, pos, b, e)) return -1;
    return parse_real(s.data() + b, s.data() + e, v) ? 0 : 1;
  };
  double energy = 0.0;
  size_t k;
  if ((k = a.find("energy=")) != std::string::npos || (k = a.find("energy:")) != std::string::npos) {
    double v = 0.0;
    if (read_first(a.substr(k + 7), &v) == 0) energy = v;
    return energy;
  }
  // the first parsable float of the line: on a conversion error drop one character and retry
  const size_t n0 = a.size();
  for (size_t it = 0; it < n0; ++it) {
    if (a.empty()) break;
    double v = 0.0;
    const int io = read_first(a, &v);
    if (io > 0) {
      a.erase(0, 1);
      const size_t b = a.find_first_not_of(' ');
      a = b == std::string::npos ? "" : a.substr(b);
      continue;
    }
    if (io == 0) energy = v;
    break;
  }
  return energy;
}

}  // namespace

extern "C" {

double crest_cuda_grepenergy(const char* line) { return line ? grepenergy(line) : 0.0; }

int crest_cuda_e2i(const char* label) { return label ? e2i(label) : 0; }

int crest_cuda_rdensemble_param(const char* fname, int* nat, int* nall) {
  if (!fname || !nat || !nall) { set_error("bad argument"); return 1; }
  *nat = 0; *nall = 0;
  crest::MappedFile f;
  if (f.open_read(fname)) return 0;                     // `inquire(exist)`: absent file = empty ensemble
  std::vector<unsigned long long> ls;
  crest::index_lines(f.data, f.size, ls);
  LineReader R(f, ls);
  std::string line;
  // pass 1: the record structure (count record, comment record, `dum` atom records)
  std::vector<long long> first;                         // first atom record of every complete structure
  std::vector<long> count;
  for (;;) {
    long dum;
    const int io = read_count(R, &dum);
    if (io < 0) break;
    if (io > 0) continue;
    if (!R.next(line)) break;                           // comment line
    if (dum < 0) dum = 0;
    const long long l0 = (long long)R.k;
    if (!R.skip((size_t)dum)) break;                    // truncated last structure: not counted
    first.push_back(l0);
    count.push_back(dum);
  }
  // pass 2: a structure counts only if every atom record reads as `str x y z`
  const size_t ns = first.size();
  std::vector<char> bad(ns, 0);
  size_t total = 0;
  long common = ns ? count[0] : 0;
  for (size_t s = 0; s < ns; ++s) { total += (size_t)count[s]; if (count[s] != common) common = -1; }
  bool done = false;
  if (common > 0 && use_device_parser(total)) {
    std::vector<unsigned> flags;
    if (crest::parse_atoms_device(f, ls, first, (int)common, nullptr, flags, 1) == 0) {
      for (size_t s = 0; s < ns; ++s)
        for (long j = 0; j < common; ++j)
          if (flags[s * (size_t)common + j] == 2) { bad[s] = 1; break; }
      done = true;
    }
  }
  if (!done)
    for (size_t s = 0; s < ns; ++s) {
      R.k = (size_t)first[s];
      for (long i = 0; i < count[s]; ++i) {
        R.next(line);
        // `read(ich,*) str,x,y,z`: a symbol item followed by three reals
        size_t pos = 0, b[4], e[4];
        int n = 0;
        while (n < 4 && next_item(line, pos, b[n], e[n])) ++n;
        double v;
        if (n < 4 || !parse_real(line.data() + b[1], line.data() + e[1], &v) ||
            !parse_real(line.data() + b[2], line.data() + e[2], &v) ||
            !parse_real(line.data() + b[3], line.data() + e[3], &v)) { bad[s] = 1; break; }
      }
    }
  for (size_t s = 0; s < ns; ++s) {
    if (bad[s]) continue;
    if (count[s] > *nat) *nat = (int)count[s];
    *nall += 1;
  }
  return 0;
}

int crest_cuda_rdensemble(const char* fname, int nat, int nall, int* at, double* xyz, double* er,
                          char* comments, int comment_len, int* nread) {
  if (!fname || nat <= 0 || nall < 0 || !at || (!xyz && nall) || comment_len < 0) {
    set_error("bad argument"); return 1;
  }
  crest::MappedFile f;
  if (f.open_read(fname)) { set_error("cannot open %s", fname); return 1; }
  std::vector<unsigned long long> ls;
  crest::index_lines(f.data, f.size, ls);
  LineReader R(f, ls);
  std::string line, sym;
  if (comments) memset(comments, ' ', (size_t)comment_len * nall);
  // pass 1: the record structure; energies and comments are taken on the way
  std::vector<long long> first;
  int i = 0;
  bool eof = false;
  while (i < nall && !eof) {
    long dum;
    int io = read_count(R, &dum);
    if (io < 0) break;
    if (io > 0) continue;
    if (dum != nat) {
      // ensemble_strucskip (strucreader.f90:483-498): read records until one starts with the
      // integer nat.  (The reference then loops over the FOREIGN count, 458-466, and mis-reads the
      // structure it just found; here that structure is read whole.)
      bool stop = false;
      for (;;) {
        io = read_count(R, &dum);
        if (io < 0) { stop = true; break; }
        if (io == 0 && dum == nat) break;
      }
      if (stop) break;
    }
    if (!R.next(line)) break;
    if (er) er[i] = grepenergy(line);
    if (comments) {
      std::string c = line;
      while (!c.empty() && c.back() == ' ') c.pop_back();
      memcpy(comments + (size_t)comment_len * i, c.data(), c.size() < (size_t)comment_len ? c.size() : (size_t)comment_len);
    }
    const long long l0 = (long long)R.k;
    if (!R.skip((size_t)nat)) { eof = true; break; }
    first.push_back(l0);
    ++i;
  }
  const size_t ns = first.size();
  // pass 2: the atom records
  auto host_record = [&](size_t s, long j, bool want_sym) -> bool {
    R.k = (size_t)first[s] + (size_t)j;
    R.next(line);
    double v[3];
    if (!coordline(line, &sym, v)) return false;
    double* dst = xyz + (s * (size_t)nat + j) * 3;
    dst[0] = v[0]; dst[1] = v[1]; dst[2] = v[2];
    if (want_sym) at[j] = e2i(sym);
    return true;
  };
  bool done = false;
  if (ns && use_device_parser(ns * (size_t)nat)) {
    std::vector<unsigned> flags;
    if (crest::parse_atoms_device(f, ls, first, nat, xyz, flags, 0) == 0) {
      done = true;
      for (size_t q = 0; q < flags.size(); ++q) {
        if (flags[q] == 0) continue;
        const size_t s = q / (size_t)nat;
        const long j = (long)(q % (size_t)nat);
        if (!host_record(s, j, false)) {                // flag 1: beyond the exact fast path; flag 2: does not scan
          set_error("%s: structure %d, atom %ld: cannot read a coordinate line", fname, (int)s + 1, j + 1);
          return 1;
        }
      }
      // the symbols: `at` holds what the LAST structure read says, like the record-by-record loop
      for (long j = 0; j < nat; ++j) host_record(ns - 1, j, true);
    }
  }
  if (!done)
    for (size_t s = 0; s < ns; ++s)
      for (long j = 0; j < nat; ++j)
        if (!host_record(s, j, true)) {
          set_error("%s: structure %d, atom %ld: cannot read a coordinate line", fname, (int)s + 1, j + 1);
          return 1;
        }
  for (size_t k = (size_t)3 * nat * ns; k < (size_t)3 * nat * nall; ++k) xyz[k] = 0.0;   // `xyz = 0.0_wp`
  if (nread) *nread = (int)ns;
  if (eof) { set_error("error while reading ensemble file."); return 1; }   // strucreader.f90:476-478
  return 0;
}

int crest_cuda_wrensemble(const char* fname, int nat, int nall, const int* at, const double* xyz,
                          const double* er, const char* comments, int comment_len) {
  if (!fname || nat <= 0 || nall < 0 || !at || (!xyz && nall)) { set_error("bad argument"); return 1; }
  FILE* f = fopen(fname, "w");
  if (!f) { set_error("cannot open %s for writing", fname); return 1; }
  std::vector<char> big(1 << 20);
  setvbuf(f, big.data(), _IOFBF, big.size());
  char sym[118 + 1][3];
  for (int z = 0; z <= 118; ++z) {
    const char* s = z >= 1 ? kPSE[z - 1] : "XX";         // i2e(...,'nc'); Z outside 1..118 -> 'XX'
    sym[z][0] = s[0]; sym[z][1] = s[1] ? s[1] : ' '; sym[z][2] = 0;
  }
  for (int i = 0; i < nall; ++i) {
    fprintf(f, "  %d\n", nat);                           // '(2x,i0)'
    if (comments) {
      int n = comment_len;
      const char* c = comments + (size_t)comment_len * i;
      while (n > 0 && (c[n - 1] == ' ' || c[n - 1] == 0)) --n;
      fwrite(c, 1, (size_t)n, f);
      fputc('\n', f);
    } else if (er) {
      fprintf(f, " %22.14f\n", er[i]);
    } else {
      fputc('\n', f);
    }
    for (int j = 0; j < nat; ++j) {
      const int z = (at[j] >= 1 && at[j] <= 118) ? at[j] : 0;
      const double* p = xyz + ((size_t)i * nat + j) * 3;
      fprintf(f, " %s %20.10f%20.10f%20.10f\n", sym[z], p[0], p[1], p[2]);   // '(1x,a2,1x,3f20.10)'
    }
  }
  const int rc = ferror(f);
  if (fclose(f) != 0 || rc) { set_error("write to %s failed", fname); return 1; }
  return 0;
}

}  // extern "C"
