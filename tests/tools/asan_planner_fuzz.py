"""Importer + planner under AddressSanitizer, no GPU needed (tfs_session_resolve plans without a device).

Build the instrumented host side next to the regular device objects and run:

    mkdir -p oracle/_build/asan
    for f in graph capi session host_staging comm; do
      nvcc -O1 -g -std=c++17 -gencode arch=compute_100a,code=sm_100a \
           -Xcompiler -fPIC,-fsanitize=address,-fno-omit-frame-pointer \
           -x cu -c tensorframes_b200/csrc/$f.cpp -o oracle/_build/asan/$f.o; done
    nvcc -shared -gencode arch=compute_100a,code=sm_100a -Xcompiler -fsanitize=address \
         -o oracle/_build/asan/libtfsb200.so oracle/_build/asan/*.o build/obj/kernels_*.o
    ASAN_OPTIONS=detect_leaks=0:halt_on_error=1:protect_shadow_gap=0 \
    LD_PRELOAD="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libstdc++.so.6)" \
        python tests/tools/asan_planner_fuzz.py 30000 7

Every mutated graph that still imports is planned for every node as a fetch, in block and in row mode, with
placeholder dimensions drawn from {0, 1, 16, 2^40}.  Any outcome but a sanitizer report is fine.
Round 1: 36000 mutations and the whole CPU suite ran clean against the instrumented build; 20000 more under
-fsanitize=undefined (same recipe, libubsan preloaded) reported nothing either.  Round 2 (comm.cpp added to the list of
instrumented files; the library now also holds tfs_graph_analyze, the pack / unpack helpers and the exchange steps):
12000 mutations and the 421 CPU tests, clean; the CPU tests once more under -fsanitize=undefined, clean.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from tensorframes_b200 import _native  # noqa: E402

ASAN_LIB = os.path.join(ROOT, "oracle", "_build", "asan", "libtfsb200.so")
if os.path.exists(ASAN_LIB):
    _native.LIB_PATH = ASAN_LIB

import _mutate  # noqa: E402


def main(n, seed):
    graphs = _mutate.seed_graphs()
    rng = np.random.default_rng(seed)
    stats = dict(imported=0, planned=0, refused=0)
    for it in range(n):
        g = _mutate.mutate(rng, graphs[it % len(graphs)], structural=bool(it & 1))
        try:
            gr = _native.Graph(g)
        except _native.NativeError:
            continue
        stats["imported"] += 1
        try:
            feeds, fetches = {}, []
            for name, _op in gr.nodes():
                dt, shape, is_ph, _ = gr.node_info(name)
                if not is_ph:
                    fetches.append(name)
                    continue
                if shape is None:
                    raise ValueError("placeholder of unknown rank")
                dims = [int(rng.choice([0, 1, 16, 1 << 40])) if d < 0 else d for d in shape.dims]
                feeds[name] = _native.DeviceArray(256, dt, tuple(dims))      # resolve never touches the data
            with _native.Session(gr, 0) as s:
                for rows in (False, True):
                    for f in fetches:
                        try:
                            s.resolve(feeds, [f], rows=rows)
                            stats["planned"] += 1
                        except _native.NativeError:
                            stats["refused"] += 1
        except (ValueError, UnicodeDecodeError, AttributeError, _native.NativeError):
            stats["refused"] += 1
        finally:
            gr.close()
    print(_native.LIB_PATH, stats)


if __name__ == "__main__":
    main(int(sys.argv[1]) if len(sys.argv) > 1 else 5000, int(sys.argv[2]) if len(sys.argv) > 2 else 0)
