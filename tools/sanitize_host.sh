#!/bin/bash
# ASan + UBSan build of the plugin host (command dispatcher + scene XML extractor) and a run of the host-side tests
# against it (parse-only modules: no GPU).  Writes profiles/r02_asan_ubsan_host.log.
set -e
cd "$(dirname "$0")/.."
C=probability-of-collision-for-safe-planning_b200/csrc
g++ -std=c++17 -O1 -g -fPIC -shared -fsanitize=address,undefined -fno-omit-frame-pointer -Wall -Wextra -Iinclude -I$C -I/usr/local/cuda/include \
    -o $C/libmcsim_asan.so $C/mc_module.cpp $C/mc_module_run.cpp $C/scene_xml.cpp $C/collada_hulls.cpp $C/mcsim_capi.cpp -L$C -lmcsim_cuda -ldl -lz -Wl,-rpath,'$ORIGIN'
ASAN="$(gcc -print-file-name=libasan.so) $(gcc -print-file-name=libstdc++.so.6)"   # libstdc++ up front: ASan must find __cxa_throw at start-up
LD_PRELOAD="$ASAN" ASAN_OPTIONS=detect_leaks=0:halt_on_error=1 UBSAN_OPTIONS=halt_on_error=1:print_stacktrace=1 python - <<'PY' 2>&1 | tee profiles/r02_asan_ubsan_host.log
import os, sys
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import pcoll_b200
pcoll_b200.plugin.LIB_PATH = os.path.join(os.path.dirname(pcoll_b200.plugin.LIB_PATH), "libmcsim_asan.so")
import pytest
rc = pytest.main(["-q", "-x", "tests/test_plugin_host.py", "tests/test_collada_hulls.py", "-p", "no:cacheprovider"])
print("ASan + UBSan build of libmcsim (g++ -fsanitize=address,undefined): tests/test_plugin_host.py + tests/test_collada_hulls.py exit code", int(rc))
sys.exit(int(rc))
PY
