#!/bin/bash
cd /root/repo
P=ctc-attention-mispronunciation_b200
mkdir -p scratch/var
python -c "import mddctc; mddctc._lib.build()" > /dev/null 2>&1 || { echo build failed; exit 1; }
n=dbg
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -Xptxas -v \
      -DMDD_PAIR_DEBUG -DMDD_PAIR_ONLY8 -c $P/csrc/ctc_pair.cu -o scratch/var/ctc_pair_$n.o > scratch/var/$n.log 2>&1 && \
    nvcc -shared -o scratch/var/libmddctc_$n.so $(ls $P/build/*.o | grep -v ctc_pair.o) scratch/var/ctc_pair_$n.o -lcudart
/usr/local/graft/bin/gpurun --timeout 600 -- "MDD_CTC_DECAY_MAX=100000 MDD_SO_PATH=scratch/var/libmddctc_$n.so python scratch/dbg_silent.py > gpurun_out/dbg_silent.txt 2>&1; echo '--- default lib'; python scratch/dbg_scale.py 2>/dev/null" 2>&1 | grep -v "^\[gpurun\] sending" | cut -c1-230
