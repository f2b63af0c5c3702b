#!/bin/bash
# same-box A/B of package builds: tools/ab.sh TAG "wl1 wl2" variant...   (variant = dir under _ab/, or "new")
cd "$(dirname "$0")/.."
T=$1; WLS=$2; shift 2
for i in 1 2; do for v in "$@"; do for wl in $WLS; do
  if [ $v == new ]; then unset SPA_PKG_DIR; else export SPA_PKG_DIR=$PWD/_ab/$v; fi
  python bench.py --workload $wl --steps 40 --no-extra --no-cpu-baseline > gpurun_out/${T}_${wl}_${v}_$i.json 2> gpurun_out/${T}_err.txt
done; done; done
python tools/benchsum.py gpurun_out/${T}_c*.json
