#!/usr/bin/env python
"""Write the binary sidecar of a local-paths text file: `<file>.epvb` next to `<file>` (2 bytes per row + 8 per
jump; epv_pathfile_write_binary).  The read_paths drop-in of the B200 build (host/read_paths_b200.cpp) then reads
the sidecar instead of parsing the text, as long as the text file is not newer.  Host only, no GPU needed.

  python tools/paths_sidecar.py sim.paths [more.paths ...]
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from epievo_b200.paths import read_paths_native, write_paths_binary   # noqa: E402


def main():
    if len(sys.argv) < 2:
        sys.exit(__doc__)
    for f in sys.argv[1:]:
        t0 = time.perf_counter()
        ps = read_paths_native(f)
        t1 = time.perf_counter()
        write_paths_binary(f + ".epvb", ps, ps.node_names)
        t2 = time.perf_counter()
        print("%s: %d sites x %d nodes, %d jumps; text %.1f MB read in %.2f s -> %s %.1f MB written in %.2f s" % (
            f, ps.n_sites, ps.n_nodes, len(ps.jumps), os.path.getsize(f) / 1e6, t1 - t0, f + ".epvb",
            os.path.getsize(f + ".epvb") / 1e6, t2 - t1))


if __name__ == "__main__":
    main()
