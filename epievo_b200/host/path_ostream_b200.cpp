/* path_ostream_b200.cpp -- drop-in for the reference's operator<<(std::ostream &, const Path &)
 * (Path.cpp:62-71): same bytes -- init_state, tot_time and every jump time in the stream's general format
 * at max_digits10 = 17 significant digits, tab-separated with a trailing tab -- formatted with
 * std::to_chars into one buffer and written with one os.write, instead of a precision change plus one
 * formatted insertion per number.  The reference's programs rewrite the whole local-paths file through this
 * operator after every MCEM iteration (prog/epievo_est_params_histories.cpp:56-75, :281-284); with the sweep
 * on the GPU that rewrite is the largest term of an iteration (tools/cli_bench.py).  Link it BEFORE the
 * reference's Path.o with -Wl,--allow-multiple-definition, like read_paths_b200.o.  Needs C++17 <charconv>. */
#include <charconv>
#include <ostream>
#include <vector>

#include "Path.hpp"

std::ostream &
operator<<(std::ostream &os, const Path &p) {
  // 24 bytes hold any double at 17 significant digits in general format, +1 for the tab
  char small[1024];
  std::vector<char> big;
  const size_t need = 2 + 25*(p.jumps.size() + 1);
  char *buf = small;
  if (need > sizeof small) { big.resize(need); buf = big.data(); }
  char *e = buf;
  *e++ = p.init_state ? '1' : '0';
  *e++ = '\t';
  e = std::to_chars(e, buf + need, p.tot_time, std::chars_format::general, 17).ptr;
  *e++ = '\t';
  for (size_t k = 0; k < p.jumps.size(); ++k) {
    e = std::to_chars(e, buf + need, p.jumps[k], std::chars_format::general, 17).ptr;
    *e++ = '\t';
  }
  os.write(buf, e - buf);
  return os;
}
