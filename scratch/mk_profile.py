"""Turn gpurun_out/prof_pair_final.ncu-rep (+ launch list, bench line) into the tracked summaries under profiles/."""
import csv, json, subprocess, shutil, sys
rep = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/prof_pair_final.ncu-rep"
subprocess.run("ncu -i %s --page raw --csv > /tmp/rawf.csv" % rep, shell=True, check=True)
subprocess.run("ncu -i %s --page source --csv --print-source cuda,sass > /tmp/srcf.csv 2>/dev/null" % rep, shell=True, check=True)
rows = list(csv.reader(open('/tmp/rawf.csv')))
hdr, units, vals = rows[0], rows[1], rows[2]
d = {h: (u, v) for h, u, v in zip(hdr, units, vals)}
keys = ['dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__time_duration.sum',
        'launch__block_size', 'launch__grid_size', 'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sector_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed']
keys += sorted(k for k in d if k.startswith('smsp__average_warps_issue_stalled') and k.endswith('per_issue_active.ratio'))
out = ["# ncu --set full --clock-control none --import-source on -k regex:ctc_pair -s 3 -c 1 python bench.py --steps 2 --warmup 3 --no-cpu",
       "# kernel: ctc_pair_kernel<NSL=8,REGBUF,255> (two utterances per warp, packed FP32), BASELINE config 3 (T=1000,B=4096,C=42,L=120), one launch",
       "# algorithmic bytes per launch 1.376e9 (4*C*sum(T_b) read + 4*C*T*B written); 4.096e6 utterance-frames"]
for k in keys:
    if k in d:
        out.append("%-100s %-16s %s" % (k, d[k][0], d[k][1]))
lines = subprocess.run("python scratch/ncu_lines.py /tmp/srcf.csv 40", shell=True, capture_output=True, text=True).stdout
out += ["", "# instructions executed / stall samples / shared-memory wavefronts (x = actual / ideal) by CUDA source line, top 40;",
        "# the lines of the inlined packed-FP32 intrinsics (sm_100_rt.hpp) also carry the emission gathers (LDS) feeding them", lines]
open('profiles/r1_ncu_full_ctc_pair.txt', 'w').write("\n".join(out) + "\n")
r, w = float(d['dram__bytes_read.sum'][1]), float(d['dram__bytes_write.sum'][1])
assert d['dram__bytes_read.sum'][0] == 'Gbyte' and d['dram__bytes_write.sum'][0] == 'Gbyte'
t = json.load(open('profiles/ncu_traffic.json'))
e = {"bytes": int((r + w) * 1e9), "kernel": "ctc_pair_kernel<8,true,255>", "dram__bytes_read.sum": int(r * 1e9),
     "dram__bytes_write.sum": int(w * 1e9), "source": "profiles/r1_ncu_full_ctc_pair.txt (ncu --set full, one launch, BASELINE config 3)"}
t['c3/algo0'] = e
t['c3/algo2'] = e
json.dump(t, open('profiles/ncu_traffic.json', 'w'), indent=1)
shutil.copy("gpurun_out/launches_pair.csv", "profiles/r1_launches_bench_pair.csv")
shutil.copy("gpurun_out/bench_pair.json", "profiles/r1_bench_pair.json")
print("time", d['gpu__time_duration.sum'], "read", r, "write", w)
