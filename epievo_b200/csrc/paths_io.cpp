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
//
// Binary sidecar ("optional binary CSR sidecar"; no counterpart in the reference): the same
// content in 2 bytes per row + 8 bytes per jump instead of ~22 bytes of text per row --
//   0   char[8] "EPVPATH1"
//   8   u64 n_sites, n_nodes, n_jumps, names_bytes
//   40  u64 checksum of everything after this 48-byte header (block hash, see payload_hash)
//   48  node names, NUL-terminated, padded to 8 bytes
//       f64 tot_time[n_nodes]
//       u16 code[n_nodes*n_sites] (bit 15 = init_state, bits 0-14 = jumps of the row; node-major), padded to 8
//       f64 jumps[n_jumps]
// little-endian, written by epv_pathfile_write_binary; epv_pathfile_read recognises the magic, so every
// consumer of the text reader (the read_paths drop-in included) reads either form.
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
  // binary sidecar: the payload as read (the file stays mapped until the handle is freed)
  bool binary = false;
  void *map = nullptr; uint64_t map_size = 0;
  const uint16_t *code = nullptr;
  const double *bin_jumps = nullptr;
  ~epv_pathfile() { if (map) munmap(map, map_size); }
};

namespace {

constexpr char BIN_MAGIC[8] = {'E', 'P', 'V', 'P', 'A', 'T', 'H', '1'};
constexpr uint64_t BIN_HEADER = 48, HASH_BLOCK = 1u << 20;

inline uint64_t rotl64(uint64_t x, int r) { return (x << r) | (x >> (64 - r)); }

// Hash of the payload: every 1 MiB block is hashed on its own (so threads can share the work and the result
// does not depend on their number), the block hashes are chained in order.
uint64_t block_hash(const unsigned char *p, uint64_t n) {
  uint64_t h = 0x9e3779b97f4a7c15ull ^ n;
  uint64_t i = 0;
  for (; i + 8 <= n; i += 8) {
    uint64_t w; std::memcpy(&w, p + i, 8);
    h = rotl64(h ^ (w*0xff51afd7ed558ccdull), 27)*0xc4ceb9fe1a85ec53ull + 0x52dce729u;
  }
  uint64_t tail = 0;
  if (i < n) std::memcpy(&tail, p + i, n - i);
  h = rotl64(h ^ (tail*0xff51afd7ed558ccdull), 27)*0xc4ceb9fe1a85ec53ull + (n - i);
  return h ^ (h >> 29);
}

uint64_t payload_hash(const unsigned char *p, uint64_t n, int n_threads) {
  const uint64_t blocks = (n + HASH_BLOCK - 1)/HASH_BLOCK;
  std::vector<uint64_t> hb(blocks);
  const int T = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)std::max(1, n_threads), blocks));
  run_threads(T, [&](int i) {
    for (uint64_t b = blocks*(uint64_t)i/T; b < blocks*(uint64_t)(i + 1)/T; ++b)
      hb[b] = block_hash(p + b*HASH_BLOCK, std::min(HASH_BLOCK, n - b*HASH_BLOCK));
  });
  uint64_t h = 0x243f6a8885a308d3ull ^ n;
  for (uint64_t b = 0; b < blocks; ++b) h = rotl64(h, 5) ^ (hb[b]*0x9e3779b97f4a7c15ull);
  return h;
}

inline uint64_t pad8(uint64_t x) { return (x + 7) & ~7ull; }

// The mapped file is a binary sidecar: validate it and point the handle at its payload.
int open_binary(epv_pathfile *f, const char *filename, void *map, uint64_t size, int n_threads) {
  const unsigned char *b = static_cast<const unsigned char *>(map);
  const std::string name(filename);
  if (size < BIN_HEADER) return io_fail("truncated binary paths file: " + name);
  uint64_t h[5];
  std::memcpy(h, b + 8, sizeof h);
  const uint64_t n_sites = h[0], n_nodes = h[1], n_jumps = h[2], names_bytes = h[3], sum = h[4];
  if (n_nodes < 2 || n_nodes > (1u << 20) || n_sites > (1ull << 40) || n_jumps > (1ull << 44) || names_bytes > (1u << 26))
    return io_fail("bad header in binary paths file: " + name);
  const uint64_t rows = n_nodes*n_sites;
  const uint64_t o_names = BIN_HEADER, o_tot = o_names + pad8(names_bytes), o_code = o_tot + 8*n_nodes,
                 o_jmp = o_code + pad8(2*rows), o_end = o_jmp + 8*n_jumps;
  if (size != o_end) return io_fail("truncated binary paths file: " + name);
  if (payload_hash(b + BIN_HEADER, size - BIN_HEADER, n_threads) != sum)
    return io_fail("checksum mismatch in binary paths file: " + name);
  const char *nm = reinterpret_cast<const char *>(b + o_names), *nm_end = nm + names_bytes;
  for (uint64_t v = 0; v < n_nodes; ++v) {
    const char *z = static_cast<const char *>(std::memchr(nm, 0, nm_end - nm));
    if (!z) return io_fail("bad name table in binary paths file: " + name);
    f->names.emplace_back(nm, z);
    nm = z + 1;
  }
  f->tot_time.resize(n_nodes);
  std::memcpy(f->tot_time.data(), b + o_tot, 8*n_nodes);
  f->code = reinterpret_cast<const uint16_t *>(b + o_code);
  f->bin_jumps = reinterpret_cast<const double *>(b + o_jmp);
  // the counts must add up (a file whose checksum matches was written by us, but the reader trusts nothing)
  const int T = pick_threads(n_threads, 2*rows);
  std::vector<uint64_t> part(T, 0);
  run_threads(T, [&](int i) {
    uint64_t acc = 0;
    for (uint64_t r = rows*(uint64_t)i/T; r < rows*(uint64_t)(i + 1)/T; ++r) acc += f->code[r] & 0x7fffu;
    part[i] = acc;
  });
  uint64_t total = 0;
  for (uint64_t x : part) total += x;
  if (total != n_jumps) return io_fail("row counts do not add up in binary paths file: " + name);
  for (uint64_t s2 = 0; s2 < n_sites; ++s2)
    if (f->code[s2]) return io_fail("the root block of a local-paths file must be empty");
  f->binary = true; f->map = map; f->map_size = size;
  f->n_sites = n_sites; f->n_nodes = n_nodes; f->n_jumps = n_jumps;
  return 0;
}

}  // namespace

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
  if (size >= 8 && std::memcmp(buf, BIN_MAGIC, 8) == 0) {          // binary sidecar
    if (open_binary(f, filename, map, size, n_threads <= 0 ? (int)std::thread::hardware_concurrency() : n_threads)) {
      f->map = nullptr; munmap(map, size); delete f; return 1;
    }
    *out = f;
    if (n_sites) *n_sites = f->n_sites;
    if (n_nodes) *n_nodes = f->n_nodes;
    if (n_jumps) *n_jumps = f->n_jumps;
    return 0;
  }
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
  if (f->binary) {
    const uint64_t rows = f->n_nodes*n;
    const int T = pick_threads(0, 4*rows);
    std::vector<uint64_t> first(T + 1, 0);             // jumps before each thread's rows
    run_threads(T, [&](int i) {
      uint64_t acc = 0;
      for (uint64_t r = rows*(uint64_t)i/T; r < rows*(uint64_t)(i + 1)/T; ++r) acc += f->code[r] & 0x7fffu;
      first[i + 1] = acc;
    });
    for (int i = 0; i < T; ++i) first[i + 1] += first[i];
    run_threads(T, [&](int i) {
      uint64_t j = first[i];
      for (uint64_t r = rows*(uint64_t)i/T; r < rows*(uint64_t)(i + 1)/T; ++r) {
        init_state[r] = (uint8_t)(f->code[r] >> 15);
        row_off[r] = j;
        j += f->code[r] & 0x7fffu;
      }
      const uint64_t j0 = first[i], j1 = first[i + 1];
      if (j1 > j0) std::memcpy(jumps + j0, f->bin_jumps + j0, (j1 - j0)*sizeof(double));
    });
    row_off[rows] = f->n_jumps;
    for (uint64_t b = 0; b < f->n_nodes; ++b) tot_time[b] = f->tot_time[b];
    return 0;
  }
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

int epv_pathfile_write_binary(const char *filename, uint64_t n_sites, uint64_t n_nodes,
                              const char *const *node_names, const uint8_t *init_state,
                              const uint64_t *row_off, const double *jumps, const double *tot_time,
                              int n_threads) {
  if (!filename || n_nodes < 2 || !node_names || !init_state || !row_off || !tot_time) return io_fail("bad argument");
  const uint64_t rows = n_nodes*n_sites, n_jumps = row_off[rows] - row_off[0];
  if (row_off[n_sites] != row_off[0]) return io_fail("the root rows of a local-paths file must be empty");
  uint64_t names_bytes = 0;
  for (uint64_t v = 0; v < n_nodes; ++v) names_bytes += std::strlen(node_names[v]) + 1;
  const uint64_t o_names = BIN_HEADER, o_tot = o_names + pad8(names_bytes), o_code = o_tot + 8*n_nodes,
                 o_jmp = o_code + pad8(2*rows), o_end = o_jmp + 8*n_jumps;
  const int fd = open(filename, O_RDWR | O_CREAT | O_TRUNC, 0644);
  if (fd < 0) return io_fail(std::string("bad output file: ") + filename);
  if (ftruncate(fd, (off_t)o_end) != 0) { close(fd); return io_fail(std::string("cannot size: ") + filename); }
  void *map = mmap(nullptr, o_end, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
  close(fd);
  if (map == MAP_FAILED) return io_fail(std::string("cannot map: ") + filename);
  unsigned char *b = static_cast<unsigned char *>(map);
  std::memset(b, 0, o_tot);                                  // header + name table + its padding
  char *nm = reinterpret_cast<char *>(b + o_names);
  for (uint64_t v = 0; v < n_nodes; ++v) { const size_t l = std::strlen(node_names[v]) + 1; std::memcpy(nm, node_names[v], l); nm += l; }
  std::memcpy(b + o_tot, tot_time, 8*n_nodes);
  const double zero = 0.0;
  std::memcpy(b + o_tot, &zero, 8);                          // the root has no branch
  uint16_t *code = reinterpret_cast<uint16_t *>(b + o_code);
  if (pad8(2*rows) != 2*rows) std::memset(b + o_code + 2*rows, 0, pad8(2*rows) - 2*rows);
  const int T = pick_threads(n_threads, 2*rows + 8*n_jumps);
  std::vector<int> too_long(T, 0);
  run_threads(T, [&](int i) {
    for (uint64_t r = rows*(uint64_t)i/T; r < rows*(uint64_t)(i + 1)/T; ++r) {
      const uint64_t c = row_off[r + 1] - row_off[r];
      if (c > 0x7fffu) { too_long[i] = 1; return; }
      code[r] = (uint16_t)(c | (init_state[r] ? 0x8000u : 0u));
    }
    const uint64_t j0 = n_jumps*(uint64_t)i/T, j1 = n_jumps*(uint64_t)(i + 1)/T;
    if (j1 > j0) std::memcpy(b + o_jmp + 8*j0, jumps + row_off[0] + j0, 8*(j1 - j0));
  });
  int rc = 0;
  for (int x : too_long) if (x) rc = io_fail("a row holds more than 32767 jumps: not representable in the binary paths format");
  for (uint64_t s2 = 0; s2 < n_sites && !rc; ++s2) if (code[s2] & 0x8000u) code[s2] = 0;     // root rows carry no state
  if (!rc) {
    std::memcpy(b, BIN_MAGIC, 8);
    const uint64_t h[5] = {n_sites, n_nodes, n_jumps, names_bytes,
                           payload_hash(b + BIN_HEADER, o_end - BIN_HEADER, T)};
    std::memcpy(b + 8, h, sizeof h);
  }
  munmap(map, o_end);
  if (rc) unlink(filename);
  return rc;
}

}  // extern "C"
