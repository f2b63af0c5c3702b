#!/bin/bash
# SASS instruction counts of the built library -> profiles/r02_sass_digest.txt (no GPU needed)
cd "$(dirname "$0")/.."
LIB=cann-ops-adv-vllm_b200/libspa_b200.so
SHA=$(python -c "import bench; print(bench.kernel_source_sha())" 2>/dev/null | tail -1)
S=$(mktemp); cuobjdump -sass $LIB > $S
NK=$(grep -c "Function : .*spa_chain_kernel" $S); NS=$(grep -c "Function : .*spa_select_kernel" $S); NA=$(grep -c "Function : .*spa_assign_kernel" $S)
{
echo "# SASS digest of $LIB (round 2; cuobjdump -sass | grep -c, whole library: $NK chain-kernel instantiations -- half of them the options build --, $NS select-kernel, $NA assign-kernel)"
echo "# kernel sources $SHA"
for m in UTMALDG UBLKCP UTMAPF SYNCS HMMA.16816 IMMA.16832 LDSM MOVM LDGSTS REDUX ATOMS UTCHMMA UTCIMMA UTCQMMA LDTM STTM HGMMA; do
  printf "%-12s %s\n" $m $(grep -c "$m" $S)
done
echo "# TMA (UTMALDG: 4-D K/V page-head slices and 2-D centroid tiles; UBLKCP: bulk copies of the staged query rows and, in the options build, of mask / pse slices), mbarrier (SYNCS),"
echo "# legacy tensor path HMMA (fp16/bf16) and IMMA (int8 logits); no tcgen05 (UTC*MMA, LDTM/STTM): see DESIGN.md section 3 for the measured reason."
} > profiles/r02_sass_digest.txt
rm -f $S; cat profiles/r02_sass_digest.txt
