"""Turn one gpurun_out/<tag>_* set (tools/gpu_round.sh) into the committed summaries under profiles/.
usage: python tools/make_profiles.py <tag> <commit>"""
import csv, json, os, subprocess, sys
tag, commit = sys.argv[1], sys.argv[2]
go, pr = "gpurun_out", "profiles"
raw = subprocess.run(["ncu", "-i", f"{go}/{tag}_step.ncu-rep", "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sectors.sum', 'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed']
kn = hdr.index('Kernel Name')
seen, traffic, ktrace = set(), {}, {}
with open(f"{pr}/{tag}_step_kernels.txt", "w") as f:
    f.write(f"# ncu --set full --clock-control none, one evaluation step of tools/dev_bench.py --gsah (config 2), commit {commit}\n")
    f.write(f"# command: tools/gpu_prof.sh {tag}; report gpurun_out/{tag}_step.ncu-rep (not committed)\n")
    for r in rows[2:]:
        name = r[kn].split('(')[0]
        if name in seen:
            continue
        seen.add(name)
        f.write(name + "\n")
        for w in want:
            if w in hdr:
                i = hdr.index(w)
                f.write('    %-72s %s %s\n' % (w, r[i], units[i]))
        def mb(w):
            i = hdr.index(w); v = float(r[i].replace(',', ''))
            return v * {'Mbyte': 1e6, 'Kbyte': 1e3, 'Gbyte': 1e9, 'byte': 1}.get(units[i], 1)
        traffic[name] = int(mb('dram__bytes_read.sum') + mb('dram__bytes_write.sum'))
        if name == 'k_trace':
            def val(w):
                return float(r[hdr.index(w)].replace(',', ''))
            ktrace = {"ms_under_ncu": val('gpu__time_duration.sum'),
                      "issue_slots_pct": val('smsp__issue_active.avg.pct_of_peak_sustained_active'),
                      "alu_pipe_pct": val('sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active'),
                      "l1_data_pipe_wavefronts_pct": val('l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed'),
                      "dram_throughput_pct": val('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed'),
                      "warp_instructions": val('smsp__inst_executed.sum'),
                      "l2_bytes": (val('lts__t_sectors.sum') * 32.0) if 'lts__t_sectors.sum' in hdr else None,
                      "l2_throughput_pct": val('lts__throughput.avg.pct_of_peak_sustained_elapsed') if 'lts__throughput.avg.pct_of_peak_sustained_elapsed' in hdr else None,
                      "registers": val('launch__registers_per_thread'),
                      "active_threads_per_instruction": val('smsp__thread_inst_executed_per_inst_executed.ratio')}
k1 = [k for k in traffic if any(x in k for x in ('view_masks', 'block_caps', 'sum_caps', 'add_u64', 'cull_emit', 'k_trace', 'k_retire'))]
json.dump({"dram_bytes_per_launch": sum(traffic[k] for k in k1),
           "source": f"ncu --set full --clock-control none, sum over the kernels of one K1 launch (config 2), profiles/{tag}_step_kernels.txt",
           "kernel_version": f"{tag} (commit {commit})", "per_kernel": {k: traffic[k] for k in k1},
           "k_trace_ncu": ktrace},
          open(f"{pr}/k1_static_r02.json", "w"), indent=1)
out = subprocess.run([sys.executable, "tools/ncu_launches.py", f"{go}/{tag}_launches.csv", "6"], capture_output=True, text=True).stdout
open(f"{pr}/{tag}_launches.txt", "w").write(
    f"# ncu --metrics gpu__time_duration.sum --clock-control none, bench.py --steps 2 --warmup 1 --no-strong --no-extras (3 resident + 3 e2e evaluations + the 24-view parity check), commit {commit}\n# K1's cub::DeviceScan::ExclusiveSum (DeviceScanInitKernel + DeviceScanKernel, ~4 us per evaluation) is outside the name filter\n" + out)
subprocess.run(["cp", f"{go}/{tag}_launches.csv", f"{pr}/{tag}_launches.csv"])
subprocess.run(["cp", f"{go}/{tag}_bench.json", f"{pr}/{tag}_bench.json"])
# k_trace source-level summary
for page, path in (("raw", "/tmp/_kt_raw.csv"), ("source", "/tmp/_kt_src.csv")):
    o = subprocess.run(["ncu", "-i", f"{go}/{tag}_step.ncu-rep", "--page", page, "--csv", "--kernel-name", "regex:^k_trace$",
                        "--launch-count", "1"], capture_output=True, text=True).stdout
    open(path, "w").write(o)
o = subprocess.run([sys.executable, "tools/ncu_summary.py", "/tmp/_kt_raw.csv", "/tmp/_kt_src.csv", "40"], capture_output=True, text=True)
open(f"{pr}/{tag}_k_trace_summary.txt", "w").write(o.stdout + o.stderr[-500:])
print(open(f"{pr}/{tag}_launches.txt").read())
print(json.load(open(f"{pr}/k1_static_r02.json"))["dram_bytes_per_launch"])
