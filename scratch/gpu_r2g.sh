timeout 600 python -m pytest tests -m gpu -x -q -k "ctc" 2>&1 | tail -4
timeout 120 python scratch/time_team.py 2>&1 | head -5
cd ctc-attention-mispronunciation_b200/csrc && nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr -DMDD_TEAM_TIMING -shared -o /tmp/libmddctc_timing.so *.cu -lcudart 2>&1 | grep -i error; cd ../..
MDD_SO_PATH=/tmp/libmddctc_timing.so python scratch/team_phases.py 2>&1 | tail -45
