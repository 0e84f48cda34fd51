// paths_io.cpp -- multi-threaded reader / writer of the reference's "local paths" text
// format, byte-compatible with
//   reader  read_paths, Path.cpp:123-148   (NODE:<name> header per node, then one line per
//           site: "<pos>\t<init>\t<tot_time>\t<jump>\t<jump>...\t")
//   writer  write_root_to_pathfile_local / append_to_pathfile_local,
//           prog/epievo_est_params_histories.cpp:56-75, with operator<<(ostream, Path),
//           Path.cpp:62-71 (doubles at max_digits10 = 17 significant digits, default float
//           format, a tab after every field).
// Host-only code: at 10^8 sites the text file is ~10 GB and the reference re-writes it
// every MCEM iteration through one iostream; here the file is cut into line-aligned chunks
// that are parsed (std::from_chars) / formatted (std::to_chars) by independent threads and
// the results land directly in the C-ABI exchange arrays (node-major CSR).
#include <algorithm>
#include <charconv>
#include <chrono>
#include <cstdlib>
#include <cstdio>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "../../include/epievo_b200.h"

namespace {

thread_local std::string t_error;

int io_fail(const std::string &m) { t_error = m; return 1; }

int pick_threads(int n, uint64_t bytes) {
  if (n <= 0) n = (int)std::thread::hardware_concurrency();
  if (n <= 0) n = 1;
  const uint64_t by_size = bytes/(1u << 20) + 1;        // at least ~1 MB of text per thread
  return (int)std::min<uint64_t>((uint64_t)std::min(n, 256), by_size);
}

template <class F>
void run_threads(int n, F f) {
  if (n == 1) { f(0); return; }
  std::vector<std::thread> th;
  for (int i = 0; i < n; ++i) th.emplace_back(f, i);
  for (auto &t : th) t.join();
}

inline bool is_node_line(const char *p, const char *e) {    // Path.cpp:52-55
  return e - p > 4 && std::memcmp(p, "NODE", 4) == 0;
}

inline const char *skip_ws(const char *p, const char *e) {
  while (p < e && (*p == ' ' || *p == '\t' || *p == '\r')) ++p;
  return p;
}

struct Chunk {
  const char *begin = nullptr, *end = nullptr;   // line-aligned
  uint64_t n_lines = 0;                          // all lines, NODE headers included
  std::vector<std::pair<uint64_t, std::string>> nodes;   // (line index within chunk, name)
  // parse results, one entry per data line in order
  std::vector<uint8_t> init;
  std::vector<uint32_t> cnt;
  std::vector<double> tot;                       // tot_time of the line
  std::vector<double> jumps;
  std::string error;
};

void scan_chunk(Chunk &c) {
  const char *p = c.begin;
  while (p < c.end) {
    const char *nl = static_cast<const char *>(std::memchr(p, '\n', c.end - p));
    const char *e = nl ? nl : c.end;
    if (is_node_line(p, e)) {
      const char *colon = static_cast<const char *>(std::memchr(p, ':', e - p));
      const char *nb = colon ? colon + 1 : p;            // get_node_name, Path.cpp:57-60
      const char *ne = e;
      if (ne > nb && ne[-1] == '\r') --ne;
      c.nodes.emplace_back(c.n_lines, std::string(nb, ne));
    }
    ++c.n_lines;
    p = nl ? nl + 1 : c.end;
  }
}

void parse_chunk(Chunk &c) {
  const uint64_t n_data = c.n_lines - c.nodes.size();
  c.init.reserve(n_data); c.cnt.reserve(n_data); c.tot.reserve(n_data);
  c.jumps.reserve((size_t)(c.end - c.begin)/24);
  const char *p = c.begin;
  while (p < c.end) {
    const char *nl = static_cast<const char *>(std::memchr(p, '\n', c.end - p));
    const char *e = nl ? nl : c.end;
    if (!is_node_line(p, e)) {
      // "<pos> <init> <tot_time> <jumps...>" (iss >> pos >> init >> tot, Path.cpp:142-145)
      const char *q = skip_ws(p, e);
      uint64_t pos = 0; unsigned init = 0; double tot = 0.0;
      auto r1 = std::from_chars(q, e, pos);
      q = skip_ws(r1.ptr, e);
      auto r2 = std::from_chars(q, e, init);
      q = skip_ws(r2.ptr, e);
      auto r3 = std::from_chars(q, e, tot);
      if (r1.ec != std::errc() || r2.ec != std::errc() || r3.ec != std::errc() || init > 1) {
        if (c.error.empty()) c.error = "malformed path line: " + std::string(p, std::min(e, p + 60));
        init = 0;
      }
      q = r3.ec == std::errc() ? r3.ptr : e;
      uint32_t n = 0;
      for (;;) {
        q = skip_ws(q, e);
        if (q >= e) break;
        double t;
        auto r = std::from_chars(q, e, t);
        if (r.ec != std::errc()) break;               // like "while (iss >> tmp_jump)"
        c.jumps.push_back(t); ++n; q = r.ptr;
      }
      c.init.push_back((uint8_t)init); c.cnt.push_back(n); c.tot.push_back(tot);
    }
    p = nl ? nl + 1 : c.end;
  }
}

}  // namespace

struct epv_pathfile {
  uint64_t n_sites = 0, n_nodes = 0, n_jumps = 0;
  std::vector<std::string> names;
  std::vector<Chunk> chunks;
  std::vector<uint64_t> chunk_row0;     // first data row (global, node-major) of each chunk
  std::vector<uint64_t> chunk_jump0;    // first jump of each chunk
  std::vector<double> tot_time;
};

extern "C" {

const char *epv_io_last_error(void) { return t_error.c_str(); }

int epv_pathfile_read(const char *filename, int n_threads, epv_pathfile **out, uint64_t *n_sites,
                      uint64_t *n_nodes, uint64_t *n_jumps) {
  if (!filename || !out) return io_fail("null argument");
  *out = nullptr;
  const int fd = open(filename, O_RDONLY);
  if (fd < 0) return io_fail(std::string("cannot read: ") + filename);     // Path.cpp:128-129
  struct stat sb;
  if (fstat(fd, &sb) != 0) { close(fd); return io_fail(std::string("cannot stat: ") + filename); }
  const uint64_t size = (uint64_t)sb.st_size;
  if (size == 0) { close(fd); return io_fail(std::string("empty paths file: ") + filename); }
  void *map = mmap(nullptr, size, PROT_READ, MAP_PRIVATE, fd, 0);
  close(fd);
  if (map == MAP_FAILED) return io_fail(std::string("cannot map: ") + filename);
  madvise(map, size, MADV_SEQUENTIAL);
  const char *buf = static_cast<const char *>(map), *end = buf + size;

  epv_pathfile *f = new epv_pathfile;
  const int T = pick_threads(n_threads, size);
  f->chunks.resize(T);
  {                                             // cut at line starts
    const char *p = buf;
    for (int i = 0; i < T; ++i) {
      f->chunks[i].begin = p;
      const char *target = i + 1 == T ? end : buf + size*(uint64_t)(i + 1)/T;
      if (target < p) target = p;
      if (target < end) {
        const char *nl = static_cast<const char *>(std::memchr(target, '\n', end - target));
        target = nl ? nl + 1 : end;
      }
      f->chunks[i].end = p = target;
    }
  }
  run_threads(T, [&](int i) { scan_chunk(f->chunks[i]); });

  // node blocks: every block after the root's must hold the same number of lines
  std::vector<uint64_t> node_line;              // global line index of every NODE header
  uint64_t line0 = 0, total_lines = 0;
  for (auto &c : f->chunks) {
    for (auto &nd : c.nodes) { node_line.push_back(line0 + nd.first); f->names.push_back(nd.second); }
    line0 += c.n_lines;
  }
  total_lines = line0;
  int rc = 0;
  f->n_nodes = node_line.size();
  if (f->n_nodes < 2 || node_line[0] != 0) rc = io_fail(std::string("not a local-paths file: ") + filename);
  if (!rc && node_line[1] != 1) rc = io_fail("the root block of a local-paths file must be empty");
  if (!rc) {
    f->n_sites = (f->n_nodes > 2 ? node_line[2] : total_lines) - node_line[1] - 1;
    for (uint64_t b = 1; b < f->n_nodes && !rc; ++b) {
      const uint64_t nxt = b + 1 < f->n_nodes ? node_line[b + 1] : total_lines;
      if (nxt - node_line[b] - 1 != f->n_sites) rc = io_fail(std::string("ragged node block in: ") + filename);
    }
  }
  if (!rc) {
    run_threads(T, [&](int i) { parse_chunk(f->chunks[i]); });
    for (auto &c : f->chunks) if (!c.error.empty() && !rc) rc = io_fail(c.error);
  }
  munmap(map, size);
  if (rc) { delete f; return rc; }

  // global positions of each chunk's rows and jumps
  f->chunk_row0.assign(T, 0); f->chunk_jump0.assign(T, 0);
  uint64_t rows = f->n_sites /* the root's (empty) rows come first */, jumps = 0;
  for (int i = 0; i < T; ++i) {
    f->chunk_row0[i] = rows; f->chunk_jump0[i] = jumps;
    rows += f->chunks[i].init.size(); jumps += f->chunks[i].jumps.size();
  }
  f->n_jumps = jumps;
  f->tot_time.assign(f->n_nodes, 0.0);
  for (int i = 0; i < T; ++i) {                 // tot_time of a node = its last row's (the rows agree)
    const Chunk &c = f->chunks[i];
    if (c.tot.empty()) continue;
    const uint64_t r0 = f->chunk_row0[i], r1 = r0 + c.tot.size() - 1;
    for (uint64_t node = r0/f->n_sites; node <= r1/f->n_sites; ++node)
      f->tot_time[node] = c.tot[std::min((node + 1)*f->n_sites - 1, r1) - r0];
  }
  *out = f;
  if (n_sites) *n_sites = f->n_sites;
  if (n_nodes) *n_nodes = f->n_nodes;
  if (n_jumps) *n_jumps = f->n_jumps;
  return 0;
}

int epv_pathfile_get(const epv_pathfile *f, uint8_t *init_state, uint64_t *row_off, double *jumps,
                     double *tot_time) {
  if (!f) return io_fail("null path file");
  const uint64_t n = f->n_sites;
  std::memset(init_state, 0, n);                       // root rows: empty (Path.hpp default)
  for (uint64_t s = 0; s <= n; ++s) row_off[s] = 0;
  const int T = (int)f->chunks.size();
  run_threads(T, [&](int i) {
    const Chunk &c = f->chunks[i];
    uint64_t r = f->chunk_row0[i], j = f->chunk_jump0[i];
    if (!c.jumps.empty()) std::memcpy(jumps + j, c.jumps.data(), c.jumps.size()*sizeof(double));
    if (!c.init.empty()) std::memcpy(init_state + r, c.init.data(), c.init.size());
    for (size_t k = 0; k < c.cnt.size(); ++k) { j += c.cnt[k]; row_off[r + k + 1] = j; }
  });
  for (uint64_t b = 0; b < f->n_nodes; ++b) tot_time[b] = f->tot_time[b];
  return 0;
}

const char *epv_pathfile_node_name(const epv_pathfile *f, uint64_t node) {
  return f && node < f->names.size() ? f->names[node].c_str() : "";
}

void epv_pathfile_free(epv_pathfile *f) { delete f; }

int epv_pathfile_write(const char *filename, uint64_t n_sites, uint64_t n_nodes,
                       const char *const *node_names, const uint8_t *init_state,
                       const uint64_t *row_off, const double *jumps, const double *tot_time,
                       int n_threads) {
  if (!filename || n_nodes < 1) return io_fail("bad argument");
  const auto t_begin = std::chrono::steady_clock::now();
  const uint64_t n_rows = (n_nodes - 1)*n_sites;
  const uint64_t est = n_rows*32 + (row_off[n_nodes*n_sites] - row_off[n_sites])*24;
  const int T = pick_threads(n_threads, est);
  std::vector<std::string> part(T);
  // thread i formats the data rows [i, i+1) * n_rows / T (node-major) and the NODE headers
  // that precede them
  run_threads(T, [&](int i) {
    std::string &o = part[i];
    const uint64_t r0 = n_rows*(uint64_t)i/T, r1 = n_rows*(uint64_t)(i + 1)/T;
    o.reserve((size_t)((r1 - r0)*40 + (row_off[n_sites + r1] - row_off[n_sites + r0])*24 + 256));
    char num[64];
    char tt[64]; size_t tt_len = 0; uint64_t tt_node = 0;
    if (i == 0) { o += "NODE:"; o += node_names[0]; o += '\n'; }
    for (uint64_t r = r0; r < r1; ++r) {
      const uint64_t node = 1 + r/n_sites, site = r % n_sites;
      if (site == 0) { o += "NODE:"; o += node_names[node]; o += '\n'; }
      if (tt_node != node) {
        tt_len = std::to_chars(tt, tt + sizeof tt, tot_time[node], std::chars_format::general, 17).ptr - tt;
        tt_node = node;
      }
      const uint64_t row = node*n_sites + site;
      char *e = std::to_chars(num, num + sizeof num, site).ptr;
      o.append(num, e); o += '\t';
      o += init_state[row] ? '1' : '0'; o += '\t';
      o.append(tt, tt_len); o += '\t';
      for (uint64_t k = row_off[row]; k < row_off[row + 1]; ++k) {
        e = std::to_chars(num, num + sizeof num, jumps[k], std::chars_format::general, 17).ptr;
        o.append(num, e); o += '\t';
      }
      o += '\n';
    }
    // nodes without any site (n_sites == 0) still get their header
    if (n_sites == 0 && i == 0)
      for (uint64_t b = 1; b < n_nodes; ++b) { o += "NODE:"; o += node_names[b]; o += '\n'; }
  });
  const bool trace = getenv("EPV_IO_TRACE") != nullptr;
  const auto t_fmt = std::chrono::steady_clock::now();
  const int fd = open(filename, O_WRONLY | O_CREAT | O_TRUNC, 0644);
  if (fd < 0) return io_fail(std::string("bad output file: ") + filename);
  std::vector<uint64_t> at(T + 1, 0);
  for (int i = 0; i < T; ++i) at[i + 1] = at[i] + part[i].size();
  int rc = 0;
  if (ftruncate(fd, (off_t)at[T]) != 0) rc = io_fail(std::string("cannot size: ") + filename);
  if (!rc) {
    std::vector<int> bad(T, 0);
    run_threads(T, [&](int i) {
      const char *p = part[i].data(); uint64_t left = part[i].size(), pos = at[i];
      while (left) {
        const ssize_t w = pwrite(fd, p, left, (off_t)pos);
        if (w <= 0) { bad[i] = 1; return; }
        p += w; left -= (uint64_t)w; pos += (uint64_t)w;
      }
    });
    for (int b : bad) if (b) rc = io_fail(std::string("write failed: ") + filename);
  }
  close(fd);
  if (trace)
    fprintf(stderr, "epv_pathfile_write: format %.3f s, write %.3f s, %d threads\n",
            std::chrono::duration<double>(t_fmt - t_begin).count(),
            std::chrono::duration<double>(std::chrono::steady_clock::now() - t_fmt).count(), T);
  return rc;
}

}  // extern "C"
