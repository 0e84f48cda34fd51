/* ostream_check.cpp -- writes 20 000 rows (sparse, dense, denormal, 18-digit values) through operator<<(ostream, Path);
 * built twice by tests/test_path_ostream.py: with the reference's Path.o alone and with host/path_ostream_b200.o in front. */
#include <fstream>
#include <iostream>
#include <random>
#include "Path.hpp"
int main(int argc, char **argv) {
  std::mt19937 g(5); std::uniform_real_distribution<double> u(0, 1);
  std::ofstream out(argv[1]);
  for (int i = 0; i < 20000; ++i) {
    Path p; p.init_state = g() & 1; p.tot_time = i % 3 == 0 ? 0.1 : u(g)*7.3;
    int n = g() % 5 == 0 ? (int)(g() % 80) : (int)(g() % 3);
    double t = 0; for (int k = 0; k < n; ++k) { t += u(g)*p.tot_time/(n + 1)*(g() % 7 == 0 ? 1e-9 : 1.0); p.jumps.push_back(t); }
    if (i % 501 == 0) p.jumps.push_back(1e-300); if (i % 777 == 0) p.jumps.push_back(123456789012345678.0);
    out << i << '\t' << p << '\n';
  }
}
