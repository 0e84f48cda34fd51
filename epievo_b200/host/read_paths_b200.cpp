/* read_paths_b200.cpp -- drop-in for the reference's read_paths (Path.cpp:123-148,
 * declared in Path.hpp) over the library's multi-threaded parser (epv_pathfile_read,
 * include/epievo_b200.h).  Same signature, same result ([node][site] Path matrix, node
 * names appended), same error ("cannot read: <file>" as std::runtime_error).  Link it
 * BEFORE the reference's Path.o with -Wl,--allow-multiple-definition (the first
 * definition wins); everything else of Path.o stays the reference's. */
#include <stdexcept>
#include <sys/stat.h>
#include <string>
#include <vector>

#include "Path.hpp"
#include "epievo_b200.h"

void
read_paths(const std::string &path_file, std::vector<std::string> &node_names,
           std::vector<std::vector<Path> > &paths) {
  epv_pathfile *f = nullptr;
  uint64_t n_sites = 0, n_nodes = 0, n_jumps = 0;
  // optional binary sidecar (epv_pathfile_write_binary, tools/paths_sidecar.py): <file>.epvb is read instead of
  // the text when it is not older than the text file -- seconds instead of minutes at 1e8 sites
  std::string source = path_file;
  {
    struct stat st_txt, st_bin;
    const std::string side = path_file + ".epvb";
    if (stat(side.c_str(), &st_bin) == 0 && stat(path_file.c_str(), &st_txt) == 0 &&
        (st_bin.st_mtim.tv_sec > st_txt.st_mtim.tv_sec ||
         (st_bin.st_mtim.tv_sec == st_txt.st_mtim.tv_sec && st_bin.st_mtim.tv_nsec >= st_txt.st_mtim.tv_nsec)))
      source = side;
  }
  if (epv_pathfile_read(source.c_str(), 0, &f, &n_sites, &n_nodes, &n_jumps)) {
    // a damaged sidecar must not hide a good text file
    if (source == path_file || epv_pathfile_read(path_file.c_str(), 0, &f, &n_sites, &n_nodes, &n_jumps))
      throw std::runtime_error(epv_io_last_error());
  }
  std::vector<uint8_t> init(n_sites*n_nodes);
  std::vector<uint64_t> off(n_sites*n_nodes + 1);
  std::vector<double> jumps(n_jumps + 1), tot(n_nodes);
  epv_pathfile_get(f, init.data(), off.data(), jumps.data(), tot.data());
  for (uint64_t b = 0; b < n_nodes; ++b) {
    node_names.push_back(epv_pathfile_node_name(f, b));
    paths.push_back(std::vector<Path>());
    if (b == 0) continue;                       // the root block of the file is empty
    std::vector<Path> &row = paths.back();
    row.resize(n_sites);
    for (uint64_t s = 0; s < n_sites; ++s) {
      const uint64_t r = b*n_sites + s;
      row[s].init_state = init[r] != 0;
      row[s].tot_time = tot[b];
      row[s].jumps.assign(jumps.begin() + off[r], jumps.begin() + off[r + 1]);
    }
  }
  epv_pathfile_free(f);
}
