#!/bin/bash
# Build the host IO library's sources together with csrc/io/fuzz_main.cc under AddressSanitizer + UBSan and run the
# mutation fuzzer; then the same program under ThreadSanitizer (loader worker threads, converter threads, threaded host
# optimizer) for a tenth of the iterations. Usage: scripts/fuzz_io.sh [iterations] [seed]
set -euo pipefail
cd "$(dirname "$0")/.."
out=$(mktemp -d)
trap 'rm -rf "$out"' EXIT
g++ -O1 -g -std=c++17 -msse4.2 -pthread -fsanitize=address,undefined -fno-sanitize-recover=undefined -fno-omit-frame-pointer \
    -Iinclude -Icsrc -o "$out/fuzz_io" csrc/io/dfm_io.cc csrc/io/tensor_bundle.cc csrc/io/host_optim.cc \
    csrc/io/libsvm_convert.cc csrc/io/fuzz_main.cc
ASAN_OPTIONS=detect_leaks=1:abort_on_error=1 UBSAN_OPTIONS=print_stacktrace=1 "$out/fuzz_io" "${1:-300}" "${2:-1}" "$out"
if [ "${DFM_FUZZ_TSAN:-1}" = "1" ]; then
  g++ -O1 -g -std=c++17 -msse4.2 -pthread -fsanitize=thread -Iinclude -Icsrc -o "$out/fuzz_tsan" csrc/io/dfm_io.cc \
      csrc/io/tensor_bundle.cc csrc/io/host_optim.cc csrc/io/libsvm_convert.cc csrc/io/fuzz_main.cc
  set +e
  TSAN_OPTIONS=halt_on_error=1 "$out/fuzz_tsan" "$(( (${1:-300} + 9) / 10 ))" "${2:-1}" "$out" > "$out/tsan.log" 2>&1
  rc=$?
  set -e
  cat "$out/tsan.log"
  if [ $rc -ne 0 ]; then
    # ThreadSanitizer cannot start under some address-space layouts (vm.mmap_rnd_bits too high): that is the tool, not a race
    if grep -q "unexpected memory mapping" "$out/tsan.log"; then
      echo "fuzz_io: ThreadSanitizer could not start on this kernel configuration; TSan pass skipped"
    else
      exit $rc
    fi
  fi
fi
