#!/bin/bash
# usage: scratch/variants.sh name1:"-DFLAG ..." name2:"..."   -> scratch/var/libmddctc_<name>.so (NSL = 8 only), then
# one gpurun call benching each at config 3 (algo 2).  Experiment tooling, not part of the product build.
cd /root/repo
P=ctc-attention-mispronunciation_b200
mkdir -p scratch/var
python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
names=()
for spec in "$@"; do
  n="${spec%%:*}"; f="${spec#*:}"; names+=("$n")
  ( nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v \
      -DMDD_PAIR_ONLY8 $f -c $P/csrc/ctc_pair.cu -o scratch/var/ctc_pair_$n.o > scratch/var/$n.log 2>&1 && \
    nvcc -shared -o scratch/var/libmddctc_$n.so $(ls $P/build/*.o | grep -v ctc_pair.o) scratch/var/ctc_pair_$n.o -lcudart ) &
done
wait
for n in "${names[@]}"; do grep -A1 "Lb1ELi255" scratch/var/$n.log | grep "spill" | sed "s/^/$n: /"; done
cmd=""
for n in "${names[@]}"; do
  cmd="$cmd MDD_SO_PATH=scratch/var/libmddctc_$n.so timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>/dev/null | python scratch/show.py '$n';"
done
/usr/local/graft/bin/gpurun --timeout 1200 -- "$cmd" 2>&1 | grep -v "^\[gpurun\] sending" | cut -c1-120
