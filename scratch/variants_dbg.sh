#!/bin/bash
# usage: scratch/variants_dbg.sh name1:"-DFLAG ..." ...  -> builds pair-kernel variants, runs scratch/dbg_scale.py + a config-3 timing for each
cd /root/repo
P=ctc-attention-mispronunciation_b200
mkdir -p scratch/var
python -c "import mddctc; mddctc._lib.build()" > /dev/null 2>&1 || { echo build failed; exit 1; }
names=()
for spec in "$@"; do
  n="${spec%%:*}"; f="${spec#*:}"; names+=("$n")
  ( nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v \
      $f -c $P/csrc/ctc_pair.cu -o scratch/var/ctc_pair_$n.o > scratch/var/$n.log 2>&1 && \
    nvcc -shared -o scratch/var/libmddctc_$n.so $(ls $P/build/*.o | grep -v ctc_pair.o) scratch/var/ctc_pair_$n.o -lcudart ) &
done
wait
for n in "${names[@]}"; do grep -A1 "Lb1ELi255ELb0" scratch/var/$n.log | grep "spill" | sed "s/^/$n: /"; done
cmd=""
for n in "${names[@]}"; do
  cmd="$cmd echo '=== $n'; MDD_SO_PATH=scratch/var/libmddctc_$n.so timeout 300 python scratch/dbg_scale.py 2>/dev/null; MDD_SO_PATH=scratch/var/libmddctc_$n.so timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>/dev/null | python scratch/show.py '$n';"
done
/usr/local/graft/bin/gpurun --timeout 1200 -- "$cmd" 2>&1 | grep -v "^\[gpurun\] sending" | cut -c1-230
