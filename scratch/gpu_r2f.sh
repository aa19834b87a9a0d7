export MDD_NVCC_EXTRA="-DMDD_TEAM_TIMING" MDD_SO_PATH=/tmp/libmddctc_timing.so
cd ctc-attention-mispronunciation_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -DMDD_TEAM_TIMING -shared -o /tmp/libmddctc_timing.so *.cu -lcudart 2>&1 | grep -i error; cd ../..
python scratch/team_phases.py 2>&1 | tail -45
