"""Summarise every kernel of an ncu --set full report: python scratch/ncu_summary.py REP OUT.txt "header line" [...]
Optionally  --traffic KEY  sums dram bytes of all kernels into profiles/ncu_traffic.json[KEY] with the source hash."""
import csv, io, json, os, subprocess, sys
sys.path.insert(0, '.')
args = sys.argv[1:]
traffic_key = None
if "--traffic" in args:
    i = args.index("--traffic")
    traffic_key = args[i + 1]
    del args[i:i + 2]
rep, out_path, headers = args[0], args[1], args[2:]
raw = open(rep).read() if rep.endswith(".csv") else subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'launch__grid_size', 'launch__block_size',
        'launch__registers_per_thread', 'launch__shared_mem_per_block_dynamic', 'launch__occupancy_limit_registers',
        'launch__occupancy_limit_shared_mem', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_st.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed']
stall = sorted(h for h in hdr if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('per_issue_active.ratio'))
ci = {h: i for i, h in enumerate(hdr)}
out = ["# " + h for h in headers]
tot_bytes = 0.0
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
names = []
for r in rows[2:]:
    if len(r) < len(hdr) - 2:
        continue
    name = r[ci["Kernel Name"]]
    names.append(name)
    out += ["", "## " + name]
    for k in keys:
        if k in ci:
            out.append("%-100s %-16s %s" % (k, units[ci[k]], r[ci[k]]))
    st = sorted(((float(r[ci[k]] or 0), k) for k in stall), reverse=True)[:6]
    for v, k in st:
        out.append("%-100s %-16s %s" % (k, units[ci[k]], v))
    for k in ('dram__bytes_read.sum', 'dram__bytes_write.sum'):
        tot_bytes += float(r[ci[k]]) * scale[units[ci[k]]]
open(out_path, "w").write("\n".join(out) + "\n")
print("wrote", out_path, len(names), "kernels, dram bytes total %.4g" % tot_bytes)
if traffic_key:
    import bench
    p = "profiles/ncu_traffic.json"
    t = json.load(open(p))
    t[traffic_key] = {"bytes": int(tot_bytes), "kernels": names, "src_sha": bench.kernel_source_hash(),
                      "source": "%s (ncu --set full, one step, all kernels of the capture summed)" % out_path}
    json.dump(t, open(p, "w"), indent=1)
    print("traffic", traffic_key, int(tot_bytes), t[traffic_key]["src_sha"])
