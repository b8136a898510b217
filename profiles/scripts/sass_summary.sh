# Opcode histogram per kernel of the shipped library: what proves the Blackwell-native paths (no GPU needed).
#   bash profiles/scripts/sass_summary.sh > profiles/r02/sass_summary.txt
cd "$(dirname "$0")/../.."
LIB=accv-lab_b200/accvlab/draw_heatmap/libdh_b200.so
echo "# cuobjdump -sass $LIB  ($(date -u +%F)), per kernel: instructions | UBLKCP (cp.async.bulk shared->global) | UGETNEXTWORKID (cluster launch control)"
echo "# | SYNCS (mbarrier) | ACQBULK+PREEXIT (griddepcontrol.wait / launch_dependents) | ATOMG/REDG (global atomics) | STG.128 / STG.other | LDG.128 / LDG.64 | MUFU.EX2 | BAR.SYNC | FADD2+FMUL2+FFMA2 (packed fp32) / FMNMX3"
cuobjdump -sass "$LIB" | awk '
/Function :/ { if (name != "") out(); name=$3; n=0; b=0; c=0; s=0; p=0; a=0; s128=0; so=0; l128=0; l64=0; ex=0; bar=0; pk=0; m3=0; next }
/^ *\/\*[0-9a-f]+\*\// {
  n++
  if ($0 ~ /UBLKCP/) b++
  if ($0 ~ /UGETNEXTWORKID/) c++
  if ($0 ~ /SYNCS/) s++
  if ($0 ~ /ACQBULK|PREEXIT/) p++
  if ($0 ~ /ATOMG|REDG|RED\.E/) a++
  if ($0 ~ /STG\.E\.128/) s128++; else if ($0 ~ /STG/) so++
  if ($0 ~ /LDG\.E\.128/) l128++; else if ($0 ~ /LDG\.E\.64/) l64++
  if ($0 ~ /MUFU\.EX2/) ex++
  if ($0 ~ /BAR\.SYNC/) bar++
  if ($0 ~ /FADD2|FMUL2|FFMA2/) pk++
  if ($0 ~ /FMNMX3/) m3++
}
function out() { printf "%-70s %6d | %3d | %2d | %3d | %2d | %3d | %3d / %3d | %3d / %3d | %3d | %2d | %3d / %3d\n", name, n, b, c, s, p, a, s128, so, l128, l64, ex, bar, pk, m3 }
END { out() }' | c++filt | sed 's/dh::RenderParams/P/; s/(bool)//g; s/(int)//g' | sort
