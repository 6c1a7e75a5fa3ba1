#!/bin/bash
# compute-sanitizer over the kernels SURVEY 5 names (D1: mbarrier pipeline without block barriers; N1: edge lists and
# fix-point state in shared memory; the append path: blocks reading one cursor word while another block writes the
# other) plus the fused head and the mAP kernels.  memcheck = global/shared out-of-bounds + misaligned + leaks of
# device-side errors; racecheck = shared-memory hazards; synccheck = barrier misuse.  Logs -> $1 (default gpurun_out/).
# usage (on a GPU box): bash tools/sanitize.sh [outdir]
OUT=${1:-gpurun_out}
mkdir -p "$OUT"
cd "$(dirname "$0")/.."
SAN=/usr/local/cuda/bin/compute-sanitizer
export YPP_SANITIZE=1
run() {   # tool, tag, command...
    tool=$1; tag=$2; shift 2
    log="$OUT/sanitizer_${tool}_${tag}.log"
    timeout 900 $SAN --tool "$tool" --print-limit 20 --error-exitcode 77 "$@" > "$log" 2>&1
    rc=$?
    echo "== $tool $tag: exit $rc; $(grep -c '=========.*\(Invalid\|hazard\|Race\|Error:\|error\)' "$log") flagged lines; $(grep 'ERROR SUMMARY\|RACECHECK SUMMARY' "$log" | tail -1)"
}
SMALL="tests/test_gpu_append.py tests/test_zz_no_alloc.py tests/test_gpu_host.py::test_batches_in_flight_append_in_batch_order tests/test_gpu_sparse_nms.py::test_sweep_and_prune_batch_of_random_segments tests/test_gpu_round2.py::test_flags_then_ap_equals_whole_matching"
run memcheck smoke python __graft_entry__.py smoke
run racecheck smoke python __graft_entry__.py smoke
run synccheck smoke python __graft_entry__.py smoke
run memcheck tests python -m pytest -x -q -m gpu $SMALL
run racecheck tests python -m pytest -x -q -m gpu $SMALL
