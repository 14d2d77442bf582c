#!/usr/bin/env bash
# Generate -> compute -> compare over the size x density grid, for one algorithm.
# The CuPy-free counterpart of the reference's cupy_cusparse/run_all_alg{1,2,3}.sh: the same
# three stages, environment knobs (SIZES, DENSITIES, CHUNK_FRACTION, PYGEN_PY, PYGEN_SCRIPT,
# CUEXE, CMPPY_PY, CMPPY_SCRIPT, STRICT), file names and [PASS]/[FAIL]/[SKIP] report lines; exit 2
# when no case was generated, and - unlike the reference, which always ends with 0 - exit 1 when
# any case FAILs.  The reference's own scripts also work unchanged with
# PYGEN_SCRIPT / CUEXE / CMPPY_SCRIPT pointed at scripts/ and spmm_b200/bin/.
#
#   usage: ALG=3 CHUNK_FRACTION=0.2 scripts/run_all_alg.sh [outdir]
set -euo pipefail

here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
root="$(dirname "$here")"
ALG="${ALG:-1}"
case "$ALG" in 1|2|3) ;; *) echo "ALG must be 1, 2 or 3" >&2; exit 2;; esac

OUTDIR="${1:-dump_alg${ALG}_txt}"
SIZES="${SIZES:-32 64 128 256 512 1024}"
DENSITIES="${DENSITIES:-0.01 0.1 0.3 0.5}"
CHUNK_FRACTION="${CHUNK_FRACTION:-0.2}"
PYGEN_PY="${PYGEN_PY:-python3}"
PYGEN_SCRIPT="${PYGEN_SCRIPT:-$here/gen_and_save_alg${ALG}_txt.py}"
CUEXE="${CUEXE:-$root/spmm_b200/bin/spgemm_from_txt_alg${ALG}}"
CMPPY_PY="${CMPPY_PY:-python3}"
CMPPY_SCRIPT="${CMPPY_SCRIPT:-$here/compare_csrs_txt.py}"
STRICT="${STRICT:-}"

gen_args=(--sizes $SIZES --densities $DENSITIES --outdir "$OUTDIR")
exe_args=()
if [ "$ALG" = 3 ]; then
  gen_args+=(--chunk-fraction "$CHUNK_FRACTION")
  exe_args+=("$CHUNK_FRACTION")
fi

echo "==[1/3] generate A/B/C(py, ALG$ALG) to $OUTDIR =="
"$PYGEN_PY" "$PYGEN_SCRIPT" "${gen_args[@]}"

echo "==[2/3] compute C(cu) with $(basename "$CUEXE") and compare =="
report="$OUTDIR/report_alg${ALG}.txt"
: > "$report"
pass=0; fail=0; total=0
shopt -s nullglob
for a_indptr in "$OUTDIR"/A_n*_dens*_alg${ALG}*_indptr.txt; do
  base="$(basename "${a_indptr%_indptr.txt}")"; base="${base#A_}"
  if [ ! -f "$OUTDIR/B_${base}_indptr.txt" ] || [ ! -f "$OUTDIR/C_py_${base}_indptr.txt" ]; then
    echo "[SKIP] loss: $base" | tee -a "$report"; continue
  fi
  "$CUEXE" "$OUTDIR/A_$base" "$OUTDIR/B_$base" "$OUTDIR/C_cu_$base" "${exe_args[@]}" >/dev/null
  if "$CMPPY_PY" "$CMPPY_SCRIPT" "$OUTDIR/C_py_$base" "$OUTDIR/C_cu_$base" $STRICT >/dev/null; then
    echo "[PASS] $base" | tee -a "$report"; pass=$((pass+1))
  else
    echo "[FAIL] $base" | tee -a "$report"; fail=$((fail+1))
  fi
  total=$((total+1))
done
if [ "$total" -eq 0 ]; then
  echo "Can't find A_* files, please check $OUTDIR content" >&2; exit 2
fi
echo "==[3/3] finish: $pass PASS / $fail FAIL / $total TOTAL =="
echo "Report: $report"
[ "$fail" -eq 0 ]
