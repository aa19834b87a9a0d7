#!/bin/bash
# build the timing variant, run the per-phase cycle counters on the GPU, restore the normal build
cd /root/repo
touch ctc-attention-mispronunciation_b200/csrc/ctc_pair.cu
MDD_NVCC_EXTRA=-DMDD_PAIR_TIMING python -c "import __graft_entry__ as g; g.build()" 2>&1 | tail -1
/usr/local/graft/bin/gpurun --timeout 600 -- "python scratch/phase_times.py > gpurun_out/$1.txt 2>&1; PT_B=296 python scratch/phase_times.py > gpurun_out/$1_b296.txt 2>&1; cat gpurun_out/$1.txt" 2>&1 | tail -16
touch ctc-attention-mispronunciation_b200/csrc/ctc_pair.cu
python -c "import __graft_entry__ as g; g.build()" 2>&1 | tail -1
