// read_paths (the reference's signature, Path.hpp) on argv[1]; prints a digest of what came back so that the
// test can compare the reference's reader on the text file with the drop-in's on text / sidecar.
#include <cstdio>
#include <iostream>
#include <string>
#include <vector>

#include "Path.hpp"

int main(int argc, char **argv) {
  if (argc < 2) return 2;
  std::vector<std::string> names;
  std::vector<std::vector<Path> > paths;
  try { read_paths(argv[1], names, paths); }
  catch (const std::exception &e) { std::cerr << e.what() << std::endl; return 1; }
  unsigned long long h = 1469598103934665603ull;
  size_t rows = 0, jumps = 0;
  auto mix = [&](unsigned long long x) { h = (h ^ x)*1099511628211ull; };
  for (size_t b = 0; b < paths.size(); ++b) {
    for (size_t c = 0; c < names[b].size(); ++c) mix((unsigned char)names[b][c]);
    mix(paths[b].size());
    for (size_t s = 0; s < paths[b].size(); ++s) {
      const Path &p = paths[b][s];
      ++rows; jumps += p.jumps.size();
      mix(p.init_state); mix(p.jumps.size());
      unsigned long long bits;
      __builtin_memcpy(&bits, &p.tot_time, 8); mix(bits);
      for (size_t j = 0; j < p.jumps.size(); ++j) { __builtin_memcpy(&bits, &p.jumps[j], 8); mix(bits); }
    }
  }
  std::printf("%zu nodes %zu rows %zu jumps digest %016llx\n", paths.size(), rows, jumps, h);
  return 0;
}
