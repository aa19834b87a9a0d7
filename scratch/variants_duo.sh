#!/bin/bash
# usage: scratch/variants_duo.sh name1:"-DFLAG ..." ...  -> scratch/var/libmddctc_<name>.so with ctc_duo.cu rebuilt under the
# flags, then one gpurun call: parity (tests -k ctc) + bench at config 3 with MDD_CTC_DUO=2 (four pairs per CTA) for each.
cd /root/repo
P=ctc-attention-mispronunciation_b200
mkdir -p scratch/var
python -c "import __graft_entry__ as g; g.build()" > /dev/null 2>&1
names=()
for spec in "$@"; do
  n="${spec%%:*}"; f="${spec#*:}"; names+=("$n")
  ( nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v \
      $f -c $P/csrc/ctc_duo.cu -o scratch/var/ctc_duo_$n.o > scratch/var/$n.log 2>&1 && \
    nvcc -shared -o scratch/var/libmddctc_$n.so $(ls $P/build/*.o | grep -v ctc_duo.o) scratch/var/ctc_duo_$n.o -lcudart ) &
done
wait
for n in "${names[@]}"; do grep -A2 "Compiling.*quad_kernel" scratch/var/$n.log | grep "spill" | sed "s/^/$n: /"; done
cmd="export MDD_CTC_DUO=2 MDD_CTC_DUO_DEBUG=1;"
for n in "${names[@]}"; do
  cmd="$cmd MDD_SO_PATH=scratch/var/libmddctc_$n.so timeout 600 python -m pytest tests/test_gpu_parity.py -q -k ctc -x 2>&1 | tail -2; MDD_SO_PATH=scratch/var/libmddctc_$n.so timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>gpurun_out/quad_$n.err | python scratch/show.py '$n'; grep ctc_quad gpurun_out/quad_$n.err;"
done
cmd="$cmd MDD_CTC_DUO=0 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu --algo 2 2>/dev/null | python scratch/show.py pair"
/usr/local/graft/bin/gpurun --timeout 1500 -- "$cmd" 2>&1 | grep -v "^\[gpurun\] sending" | cut -c1-200
