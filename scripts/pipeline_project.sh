#!/bin/bash
# pipeline_project.sh — `rake pipeline:project` (Rakefile.rb:156-215) with hovl_cli in the overlapper slot:
#   toAmos_new -s seq -b bnk ; OVERLAPPER -i seq -o ovl ; bank-transact -b bnk -m ovl ; tigger -b bnk ;
#   make-consensus -e 0.04 -o 40 -B -b bnk ; bank2fasta -b bnk > fasta          (Rakefile.rb:181-190)
# usage: scripts/pipeline_project.sh <reads.seq> <amos-bin-dir> [work-dir] [-- hovl_cli flags, e.g. -k 16]
#   HOVL_OVERLAPPER : the overlapper command (default sequence-aligner_b200/hovl_cli; the Scala host would be
#                     "scala -cp build:lib/trove.jar Project4Native")
#   HOVL_OVL        : take this .ovl instead of running the overlapper (boxes without a GPU)
# Prints the per-stage wall-clock table of the Rakefile and leaves <work-dir>/<name>.fasta.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
[ $# -ge 2 ] || { echo "usage: $0 <reads.seq> <amos-bin-dir> [work-dir] [-- overlapper flags]" >&2; exit 2; }
SEQ_IN="$1"; BF="$2"; shift 2
TF="pipeline_tmp"
if [ $# -ge 1 ] && [ "$1" != "--" ]; then TF="$1"; shift; fi
if [ $# -ge 1 ] && [ "$1" = "--" ]; then shift; fi
[ -f "$SEQ_IN" ] || { echo "No $SEQ_IN exists." >&2; exit 1; }
for t in toAmos_new bank-transact tigger make-consensus bank2fasta; do
  [ -x "$BF/$t" ] || { echo "AMOS tool $BF/$t not found" >&2; exit 3; }
done
OVERLAPPER="${HOVL_OVERLAPPER:-$HERE/sequence-aligner_b200/hovl_cli}"
rm -rf "$TF"; mkdir -p "$TF"
raw="$(basename "${SEQ_IN%.*}")"
seq="$TF/$raw.seq"; ovl="$TF/$raw.ovl"; bnk="$TF/$raw.bnk"; fst="$TF/$raw.fasta"
cp "$SEQ_IN" "$seq"
now() { date +%s.%N; }
t0=$(now)
"$BF/toAmos_new" -s "$seq" -b "$bnk" > "$TF/toAmos.log" 2>&1
t1=$(now)
if [ -n "${HOVL_OVL:-}" ]; then cp "$HOVL_OVL" "$ovl"; else $OVERLAPPER -i "$seq" -o "$ovl" "$@"; fi
t2=$(now)
"$BF/bank-transact" -b "$bnk" -m "$ovl" > "$TF/transact.log" 2>&1
t3=$(now)
"$BF/tigger" -b "$bnk" > "$TF/tigger.log" 2>&1
t4=$(now)
"$BF/make-consensus" -e 0.04 -o 40 -B -b "$bnk" > "$TF/consensus.log" 2>&1
t5=$(now)
"$BF/bank2fasta" -b "$bnk" > "$fst" 2> "$TF/bank2fasta.log"
t6=$(now)
d() { awk -v a="$1" -v b="$2" 'BEGIN { printf "%.3f s", b - a }'; }
echo ""
echo "============ Time Taken ============="
echo "Total Time               : $(d "$t0" "$t6")"
echo "  Bank Creation Time     : $(d "$t0" "$t1")"
echo "  Overlap Time           : $(d "$t1" "$t2")"
echo "  Bank Transaction Time  : $(d "$t2" "$t3")"
echo "  Contigger Time         : $(d "$t3" "$t4")"
echo "  Consensus Time         : $(d "$t4" "$t5")"
echo "  Fasta Creation Time    : $(d "$t5" "$t6")"
echo ""
echo "contigs: $(grep -c '^>' "$fst" || true)  ->  $fst"
