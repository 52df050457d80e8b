"""Turn the captures of tools/gpu_ncu.sh (gpurun_out/prof_*.ncu-rep, launches.csv, b_bf16.log) into the tracked summaries under
profiles/: usage  python tools/ncu_summarize.py <tag> "<one-line description of the build>"."""
import csv, json, os, shutil, subprocess, sys
tag, desc = sys.argv[1], sys.argv[2]
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
out_dir, go = os.path.join(root, "profiles"), os.path.join(root, "gpurun_out")
for rep, raw in (("prof_lstm.ncu-rep", "lstm_raw.csv"), ("prof_gemm.ncu-rep", "gemm_raw.csv")):
    with open(os.path.join(go, raw), "w") as f:
        subprocess.run(["ncu", "-i", os.path.join(go, rep), "--page", "raw", "--csv"], stdout=f, stderr=subprocess.DEVNULL, check=True)
shutil.copy(os.path.join(go, "launches.csv"), os.path.join(out_dir, f"r2_ncu_launches_{tag}.csv"))
shutil.copy(os.path.join(go, "b_bf16.log"), os.path.join(out_dir, f"r2_bench_{tag}.json"))
keys = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__registers_per_thread', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__m_xbar2l1tex_read_bytes.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_issued.avg',
        'smsp__average_warp_latency_per_inst_issued.ratio', 'l1tex__data_pipe_lsu_wavefronts_mem_shared_op_ld.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum']
out = [f"# ncu --set full summaries, round 2, capture {tag} (B200, config 2: B=256, T=128, bf16 mode)\n",
       f"Build: {desc}. Captured with `tools/gpu_ncu.sh`, summarised by `tools/ncu_summarize.py`. Per-launch values, cold cache,",
       "serialised - compare shares, not absolutes. Under ncu the cooperative cluster launch of the recurrence falls back to the checked",
       f"plain cluster launch (`launch.h::launch_persistent`): the identical grid. Launch list of the same command: `r2_ncu_launches_{tag}.csv`;",
       f"bench line of the same build and box: `r2_bench_{tag}.json`.\n"]
traffic = {}
for f, title in [('lstm_raw.csv', 'll_lstm_kernel (lstm1 H=512, then the two lstm2 layers H=1024)'),
                 ('gemm_raw.csv', 'tc_gemm_kernel (the 18 launches of one forward, in order)')]:
    rows = list(csv.reader(open(os.path.join(go, f)))); hdr, units = rows[0], rows[1]
    out.append(f"## {title}\n")
    out.append("| metric | unit | " + " | ".join(str(i + 1) for i in range(len(rows) - 2)) + " |")
    out.append("|---|---|" + "---|" * (len(rows) - 2))
    for k in ["Kernel Name", "Grid Size"] + keys:
        if k not in hdr: continue
        i = hdr.index(k); vals = []
        for r in rows[2:]:
            v = r[i]
            if k == "Kernel Name": v = v.replace("void unnamed>::", "").split("(")[0]
            else:
                try: v = f"{float(v.replace(',', '')):.4g}"
                except ValueError: pass
            vals.append(v)
        out.append(f"| {k} | {units[i]} | " + " | ".join(vals) + " |")
    out.append("")
    if 'lstm' in f:
        ir, iw = hdr.index('dram__bytes_read.sum'), hdr.index('dram__bytes_write.sum')
        mult = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
        per = [float(r[ir].replace(',', '')) * mult[units[ir]] + float(r[iw].replace(',', '')) * mult[units[iw]] for r in rows[2:]]
        traffic = {'lstm1': per[0], 'lstm2': per[1] + per[2]}
json.dump({"source": f"profiles/r2_ncu_full_summary_{tag}.md (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, summed over the launches of the phase)",
           "config": "B=256,T=128,bf16", "bytes_per_step": traffic}, open(os.path.join(out_dir, "r2_ncu_traffic.json"), "w"), indent=1)
open(os.path.join(out_dir, f"r2_ncu_full_summary_{tag}.md"), "w").write("\n".join(out))
print("traffic", traffic)
