// sysio.cpp -- reading and writing linear systems (host only).
//
// The reference reads its input with read_input_file (cg.c:146-218): a 4-line comma-separated
// text file -- line 0 column indices, 1 row pointers, 2 matrix values, 3 right-hand side -- taken
// apart one fgetc at a time into a 64-byte token buffer (cg.c:317-356) and appended to arrays that
// start at the size of the course's matrix and grow by 1000 entries per realloc
// (cg.c:235-243,265-273,293-302).  n is the number of row pointers minus one (cg.c:204); integers
// go through strtol, reals through strtod (cg.c:348-356).  The return code is ignored by its
// caller (cg.c:65) and the reader never terminates on a file without a final newline.
//
// Here the file is mapped, the four lines are located with memchr, every line is cut at commas
// into one piece per thread, the pieces' tokens are counted, and then parsed in parallel straight
// into malloc'ed arrays of the exact size (the caller owns them and may free() them, as it may
// the reference's).  Integers are parsed by hand; reals by std::from_chars with strtod as the
// fallback for spellings from_chars refuses (leading '+', hex floats, "inf") -- both round
// correctly, so every value has the bits the reference's strtod gives it.  Differences, all
// deliberate: errors are reported (-1 and cgb200_last_error()); a missing final newline, CR LF,
// blanks around tokens and tokens of any reasonable length are accepted; an empty token between
// two commas is an error (the reference stores strtol("") = 0 for it).
//
// A binary sidecar ("CGB200S1": n, nnz, then row_ptr, col_indices, values, b as raw little-endian
// arrays) holds the same four arrays for systems whose text form would take minutes to format and
// parse; cgb200_read_system recognises it by its magic.
#include <errno.h>
#include <fcntl.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <charconv>
#include <string>
#include <thread>
#include <vector>

#include "../../include/cgb200.h"

namespace cgb {
void set_error(const std::string& m);  // solver.cu (not included here: this file needs no CUDA)
}

namespace {

using cgb::set_error;

const char MAGIC[8] = {'C', 'G', 'B', '2', '0', '0', 'S', '1'};

int io_threads(size_t bytes) {
  const char* e = getenv("CGB200_IO_THREADS");
  int t = e ? atoi(e) : 0;
  if (t <= 0) {
    t = (int)std::thread::hardware_concurrency();
    if (t <= 0) t = 4;
    if (t > 32) t = 32;
  }
  const size_t by_size = bytes / (256u << 10) + 1;  // a thread per 256 KiB at least
  if ((size_t)t > by_size) t = (int)by_size;
  return t < 1 ? 1 : t;
}

inline bool blank(char c) { return c == ' ' || c == '\t' || c == '\r'; }

struct Piece {
  const char* b;
  const char* e;   // [b, e): whole tokens, cut at commas
  size_t ntok = 0; // tokens in the piece
  size_t first = 0;
  bool bad = false;
};

// Tokens of [b, e): fields between commas.  An empty last field is a trailing comma (accepted,
// as the reference does: nothing is pending when the newline arrives, cg.c:176-201).
size_t count_tokens(const char* b, const char* e, bool last_piece) {
  size_t commas = 0;
  for (const char* p = b; p < e; ++p) commas += (*p == ',');
  if (!last_piece) return commas;  // every field of an inner piece ends with its comma
  const char* q = e;
  while (q > b && blank(q[-1])) --q;
  const bool open_field = q > b && q[-1] != ',';
  return commas + (open_field ? 1 : 0);
}

bool parse_int_tok(const char* b, const char* e, int* out) {
  while (b < e && blank(*b)) ++b;
  while (e > b && blank(e[-1])) --e;
  if (b == e) return false;
  bool neg = false;
  if (*b == '+' || *b == '-') neg = *b++ == '-';
  if (b == e) return false;
  long long v = 0;
  if (*b < '0' || *b > '9') return false;
  for (; b < e && *b >= '0' && *b <= '9'; ++b) {
    v = v * 10 + (*b - '0');
    if (v > 0x7fffffffLL + 1) return false;
  }
  if (b < e && *b == '.')  // "3.0", as some exporters write indices; strtol reads the 3 (cg.c:348)
    for (++b; b < e && *b == '0'; ++b) {
    }
  if (b != e) return false;
  if (neg) v = -v;
  if (v > 0x7fffffffLL || v < -0x80000000LL) return false;
  *out = (int)v;
  return true;
}

bool parse_real_tok(const char* b, const char* e, double* out) {
  while (b < e && blank(*b)) ++b;
  while (e > b && blank(e[-1])) --e;
  if (b == e) return false;
  auto r = std::from_chars(b, e, *out);
  if (r.ec == std::errc() && r.ptr == e) return true;
  // spellings strtod takes and from_chars does not; also its over/underflow results
  char buf[512];
  const size_t len = (size_t)(e - b);
  if (len >= sizeof buf) return false;
  memcpy(buf, b, len);
  buf[len] = 0;
  char* end = nullptr;
  *out = strtod(buf, &end);
  return end == buf + len;
}

template <class T, class F>
void parse_piece(Piece& pc, T* dst, F parse_one) {
  const char* p = pc.b;
  size_t k = pc.first;
  const size_t stop = pc.first + pc.ntok;
  while (k < stop) {
    const char* q = (const char*)memchr(p, ',', (size_t)(pc.e - p));
    const char* te = q ? q : pc.e;
    if (!parse_one(p, te, &dst[k])) {
      pc.bad = true;
      return;
    }
    ++k;
    p = q ? q + 1 : pc.e;
  }
  // nothing but blanks may follow the last token of the line (a trailing comma is consumed above)
  for (; p < pc.e; ++p)
    if (!blank(*p) && *p != ',') {
      pc.bad = true;
      return;
    }
}

// Parses one line [b, e) into a malloc'ed array; *count receives the number of entries.
template <class T, class F>
T* parse_line(const char* b, const char* e, int nthreads, size_t* count, F parse_one, bool* bad) {
  std::vector<Piece> pieces;
  const size_t len = (size_t)(e - b);
  int np = nthreads;
  if (len < (size_t)np * 4096) np = (int)(len / 4096) + 1;
  const char* cur = b;
  for (int i = 0; i < np && cur < e; ++i) {
    const char* target = i + 1 == np ? e : b + len * (size_t)(i + 1) / (size_t)np;
    if (target < cur) target = cur;
    const char* cut = e;
    if (target < e) {
      const char* c = (const char*)memchr(target, ',', (size_t)(e - target));
      cut = c ? c + 1 : e;
    }
    Piece pc;
    pc.b = cur;
    pc.e = cut;
    pieces.push_back(pc);
    cur = cut;
  }
  if (pieces.empty()) {
    Piece pc;
    pc.b = b;
    pc.e = e;
    pieces.push_back(pc);
  }
  {
    std::vector<std::thread> th;
    for (size_t i = 0; i < pieces.size(); ++i)
      th.emplace_back([&, i] { pieces[i].ntok = count_tokens(pieces[i].b, pieces[i].e, i + 1 == pieces.size()); });
    for (auto& t : th) t.join();
  }
  size_t total = 0;
  for (auto& pc : pieces) {
    pc.first = total;
    total += pc.ntok;
  }
  T* dst = (T*)malloc(sizeof(T) * (total > 0 ? total : 1));
  if (!dst) {
    *bad = true;
    return nullptr;
  }
  {
    std::vector<std::thread> th;
    for (size_t i = 0; i < pieces.size(); ++i)
      th.emplace_back([&, i] { parse_piece(pieces[i], dst, parse_one); });
    for (auto& t : th) t.join();
  }
  for (auto& pc : pieces) *bad |= pc.bad;
  *count = total;
  return dst;
}

struct Mapped {
  const char* p = nullptr;
  size_t n = 0;
  int fd = -1;
  ~Mapped() {
    if (p && n) munmap((void*)p, n);
    if (fd >= 0) close(fd);
  }
};

int fail(const std::string& m) {
  set_error(m);
  return -1;
}

int fill_structs(struct __mv_sparse* A, struct __mv_sparse* b, const char* path, int* rp, size_t nrp,
                 int* cols, size_t ncols, double* vals, size_t nvals, double* rhs, size_t nrhs) {
  const bool ok = nrp >= 1 && nrp <= 0x7fffffffu && ncols == nvals && nrhs == nrp - 1 && rp[0] == 0 &&
                  (size_t)(unsigned)rp[nrp - 1] == nvals && nvals <= 0x7fffffffu;
  if (!ok) {
    free(rp);
    free(cols);
    free(vals);
    free(rhs);
    char m[384];
    snprintf(m, sizeof m,
             "malformed input (%s): %zu column indices, %zu row pointers, %zu values, %zu rhs entries",
             path, ncols, nrp, nvals, nrhs);
    return fail(m);
  }
  A->size = (int)nrp - 1;  // cg.c:204
  A->nnz = (int)nvals;
  A->values = vals;
  A->col_indices = cols;
  A->row_ptr = rp;
  b->size = (int)nrhs;  // cg.c:210-212
  b->nnz = (int)nrhs;
  b->values = rhs;
  b->col_indices = NULL;
  b->row_ptr = NULL;
  return 0;
}

int read_binary(const Mapped& f, const char* path, struct __mv_sparse* A, struct __mv_sparse* b) {
  if (f.n < 24) return fail(std::string("truncated binary system file: ") + path);
  int64_t n, nnz;
  memcpy(&n, f.p + 8, 8);
  memcpy(&nnz, f.p + 16, 8);
  if (n < 0 || nnz < 0 || n > 0x7ffffffe || nnz > 0x7fffffff)
    return fail(std::string("binary system file with impossible sizes: ") + path);
  const size_t rp_b = 4 * ((size_t)n + 1), co_b = 4 * (size_t)nnz;
  const size_t pad = (8 - (rp_b + co_b) % 8) % 8;
  const size_t need = 24 + rp_b + co_b + pad + 8 * (size_t)nnz + 8 * (size_t)n;
  if (f.n < need) return fail(std::string("truncated binary system file: ") + path);
  int* rp = (int*)malloc(rp_b);
  int* cols = (int*)malloc(co_b ? co_b : 4);
  double* vals = (double*)malloc(nnz ? 8 * (size_t)nnz : 8);
  double* rhs = (double*)malloc(n ? 8 * (size_t)n : 8);
  if (!rp || !cols || !vals || !rhs) {
    free(rp); free(cols); free(vals); free(rhs);
    return fail("host allocation failed");
  }
  const char* q = f.p + 24;
  // the four copies are independent and memory-bound: one thread each
  std::thread t0([&] { memcpy(rp, q, rp_b); });
  std::thread t1([&] { memcpy(cols, q + rp_b, co_b); });
  std::thread t2([&] { memcpy(vals, q + rp_b + co_b + pad, 8 * (size_t)nnz); });
  memcpy(rhs, q + rp_b + co_b + pad + 8 * (size_t)nnz, 8 * (size_t)n);
  t0.join(); t1.join(); t2.join();
  return fill_structs(A, b, path, rp, (size_t)n + 1, cols, (size_t)nnz, vals, (size_t)nnz, rhs, (size_t)n);
}

bool is_vector_struct(const struct __mv_sparse* v) { return v && !v->col_indices && !v->row_ptr; }

int check_writable(const struct __mv_sparse* A, const struct __mv_sparse* b) {
  if (!A || !b || !A->row_ptr || !A->values || (!A->col_indices && A->nnz > 0) || !is_vector_struct(b) ||
      (!b->values && b->size > 0) || A->size != b->size || A->size < 0 || A->nnz < 0)
    return fail("cgb200_write_system: A must be a CSR matrix and b a vector of the same size");
  return 0;
}

// decimal text of one piece of an array, comma separated (no trailing comma)
template <class T>
void format_piece(const T* v, size_t lo, size_t hi, std::string* out) {
  char buf[40];
  out->reserve((hi - lo) * (sizeof(T) == 8 ? 20 : 8));
  for (size_t i = lo; i < hi; ++i) {
    int len;
    if constexpr (sizeof(T) == 8) {
      // shortest text that reads back to the same bits (what "%.17g" guarantees, usually shorter)
      auto r = std::to_chars(buf, buf + sizeof buf, (double)v[i]);
      len = (int)(r.ptr - buf);
    } else {
      auto r = std::to_chars(buf, buf + sizeof buf, (int)v[i]);
      len = (int)(r.ptr - buf);
    }
    out->append(buf, (size_t)len);
    if (i + 1 < hi) out->push_back(',');
  }
}

template <class T>
bool write_line(FILE* fp, const T* v, size_t n) {
  const int nt = io_threads(n * 8);
  std::vector<std::string> parts((size_t)nt);
  std::vector<std::thread> th;
  for (int t = 0; t < nt; ++t)
    th.emplace_back([&, t] { format_piece(v, n * (size_t)t / nt, n * (size_t)(t + 1) / nt, &parts[(size_t)t]); });
  for (auto& t : th) t.join();
  bool first = true;
  for (auto& s : parts) {
    if (s.empty()) continue;
    if (!first && fputc(',', fp) == EOF) return false;
    if (fwrite(s.data(), 1, s.size(), fp) != s.size()) return false;
    first = false;
  }
  return fputc('\n', fp) != EOF;
}

}  // namespace

extern "C" {

int cgb200_read_system(const char* path, struct __mv_sparse* A, struct __mv_sparse* b) {
  if (!path || !A || !b) return fail("cgb200_read_system: NULL argument");
  Mapped f;
  f.fd = open(path, O_RDONLY);
  if (f.fd < 0) return fail(std::string("Failed to open input file (") + path + "): " + strerror(errno));
  struct stat st;
  if (fstat(f.fd, &st) != 0) return fail(std::string("fstat(") + path + "): " + strerror(errno));
  f.n = (size_t)st.st_size;
  if (f.n == 0) return fail(std::string("empty input file: ") + path);
  void* m = mmap(nullptr, f.n, PROT_READ, MAP_PRIVATE, f.fd, 0);
  if (m == MAP_FAILED) {
    f.n = 0;
    return fail(std::string("mmap(") + path + "): " + strerror(errno));
  }
  f.p = (const char*)m;
  madvise(m, f.n, MADV_SEQUENTIAL);
  if (f.n >= 8 && memcmp(f.p, MAGIC, 8) == 0) return read_binary(f, path, A, b);

  // the four lines; whatever follows them is ignored, as by the reference (cg.c:176-186)
  const char* lb[4];
  const char* le[4];
  const char* cur = f.p;
  const char* end = f.p + f.n;
  int nlines = 0;
  for (; nlines < 4 && cur <= end; ++nlines) {
    if (cur == end && nlines > 0) break;  // the previous line ended the file with its newline
    const char* nl = (const char*)memchr(cur, '\n', (size_t)(end - cur));
    lb[nlines] = cur;
    le[nlines] = nl ? nl : end;
    cur = nl ? nl + 1 : end + 1;
  }
  if (nlines < 4) {
    char msg[256];
    snprintf(msg, sizeof msg, "malformed input (%s): %d line(s), 4 expected", path, nlines);
    return fail(msg);
  }
  const int nt = io_threads(f.n);
  bool bad = false;
  size_t ncols = 0, nrp = 0, nvals = 0, nrhs = 0;
  int* cols = parse_line<int>(lb[0], le[0], nt, &ncols, parse_int_tok, &bad);
  int* rp = parse_line<int>(lb[1], le[1], nt, &nrp, parse_int_tok, &bad);
  double* vals = parse_line<double>(lb[2], le[2], nt, &nvals, parse_real_tok, &bad);
  double* rhs = parse_line<double>(lb[3], le[3], nt, &nrhs, parse_real_tok, &bad);
  if (bad || !cols || !rp || !vals || !rhs) {
    free(cols); free(rp); free(vals); free(rhs);
    return fail(std::string("malformed input (") + path + "): empty or non-numeric token");
  }
  return fill_structs(A, b, path, rp, nrp, cols, ncols, vals, nvals, rhs, nrhs);
}

int cgb200_write_system(const char* path, const struct __mv_sparse* A, const struct __mv_sparse* b,
                        int binary) {
  if (!path) return fail("cgb200_write_system: NULL path");
  if (check_writable(A, b)) return -1;
  FILE* fp = fopen(path, "wb");
  if (!fp) return fail(std::string("cannot create ") + path + ": " + strerror(errno));
  const size_t n = (size_t)A->size, nnz = (size_t)A->nnz;
  bool ok = true;
  if (binary) {
    const int64_t hdr[2] = {(int64_t)n, (int64_t)nnz};
    const size_t rp_b = 4 * (n + 1), co_b = 4 * nnz;
    const size_t pad = (8 - (rp_b + co_b) % 8) % 8;
    const char zeros[8] = {0};
    ok = fwrite(MAGIC, 1, 8, fp) == 8 && fwrite(hdr, 8, 2, fp) == 2 &&
         fwrite(A->row_ptr, 4, n + 1, fp) == n + 1 && fwrite(A->col_indices, 4, nnz, fp) == nnz &&
         fwrite(zeros, 1, pad, fp) == pad && fwrite(A->values, 8, nnz, fp) == nnz &&
         fwrite(b->values, 8, n, fp) == n;
  } else {
    ok = write_line(fp, A->col_indices, nnz) && write_line(fp, A->row_ptr, n + 1) &&
         write_line(fp, A->values, nnz) && write_line(fp, b->values, n);
  }
  if (fclose(fp) != 0) ok = false;
  if (!ok) return fail(std::string("write failed: ") + path + ": " + strerror(errno));
  return 0;
}

}  // extern "C"
