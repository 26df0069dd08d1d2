#!/usr/bin/env bash
# compute-sanitizer over small GPU cases that reach every kernel family of the library ("Race
# detection / sanitizers"). Run on the GPU box:   bash tools/run_sanitizers.sh [outdir]
# memcheck runs a pytest subset (default path, GS_SMALL_SORT both ways, generic binning through the switch test, split SH,
# exchange kernels on one device, relocation, add-noise, loss, Adam, activations, swin, knn, the no-alloc L1);
# racecheck / initcheck / synccheck run the smaller smoke() + a handful of tests (they are 10-100x slower).
# Every tool's report ends with "ERROR SUMMARY: N errors"; the script prints those lines and never fails the shell.
set -uo pipefail
cd "$(dirname "$0")/.."
OUT=${1:-gpurun_out}
mkdir -p "$OUT"
SAN="compute-sanitizer --print-limit 30 --error-exitcode 0 --target-processes all"
SMALL='test_against_golden_vectors or test_all_culled or test_empty_input or test_single_isotropic or test_duplicate_gaussians or test_saturating_stack or test_add_noise_vs_oracle or test_compute_relocation_bit_exact or test_autograd_api or test_split_sh_through_the_public or test_dp_entry_points or test_slab_all_reduce_single or test_fused_adam_state_surgery or test_fused_activations_without or test_swin_kernels_against_reference_golden or test_static_l1 or test_fused_loss_against_reference_golden or test_knn_vs_cpu_oracle or test_integration_md'
WIDE="$SMALL or test_switchable_code_paths or test_gradient_records_equal or test_factored_exchange_world_1 or test_lower_sh_degrees or test_precomputed_colors or test_unaligned_rotations or test_fused_l1_loss_vs_oracle"
run() {  # name, tool args..., -- command
    local name=$1; shift
    echo "== $name"
    timeout 1500 "$@" > "$OUT/r3_sanitizer_$name.log" 2>&1
    echo "   exit $? ; $(grep -c 'ERROR SUMMARY' "$OUT/r3_sanitizer_$name.log") summaries: $(grep 'ERROR SUMMARY' "$OUT/r3_sanitizer_$name.log" | sort | uniq -c | tr '\n' ';')"
    tail -3 "$OUT/r3_sanitizer_$name.log" | cut -c1-200
}
SMOKE='import __graft_entry__ as g; g.smoke()'
run memcheck_pytest   $SAN --tool memcheck  python -m pytest tests -m gpu -q -x -k "$WIDE"
run memcheck_smallsort0 env GS_SMALL_SORT=0 $SAN --tool memcheck python -c "$SMOKE"
run racecheck_smoke   $SAN --tool racecheck python -c "$SMOKE"
run racecheck_pytest  $SAN --tool racecheck python -m pytest tests -m gpu -q -x -k "$SMALL"
run synccheck_smoke   $SAN --tool synccheck python -c "$SMOKE"
run initcheck_smoke   $SAN --tool initcheck python -c "$SMOKE"
run initcheck_pytest  $SAN --tool initcheck python -m pytest tests -m gpu -q -x -k "$SMALL"
