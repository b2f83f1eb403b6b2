"""Reads an .ncu-rep here (no GPU): key raw metrics + the SASS lines with the most stall samples."""
import csv, subprocess, sys, io
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); hdr, units = rows[0], rows[1]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed', 'launch__registers_per_thread', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__m_xbar2l1tex_read_bytes.sum', 'l1tex__m_l1tex2xbar_write_bytes.sum', 'l1tex__t_requests_pipe_lsu_mem_local_op_st.sum',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__cycles_elapsed.max', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed']
want += [h for h in hdr if ('bank_conflict' in h or 'data_pipe_lsu_wavefronts_mem_shared' in h or 'l1tex__data_pipe_lsu_wavefronts' in h) and h not in want]
want += [h for h in hdr if 'pcsamp_warps_issue_stalled' in h and 'not_issued' not in h]
for w in want:
    if w in hdr:
        i = hdr.index(w); print(f"{w} [{units[i]}]: {[r[i] for r in rows[2:]]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]; isrc = h.index("Source"); isamp = h.index("# Samples"); iex = h.index("Instructions Executed")
stalls = [c for c in h if c.startswith("stall_") and "Not Issued" not in c]
out = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name": break
    if len(r) > isamp: out.append(r)
tot = sum(int(r[isamp]) for r in out)
print("instructions", len(out), "samples", tot)
top = sorted(range(len(out)), key=lambda i: -int(out[i][isamp]))[:ntop]
for i in sorted(top):
    r = out[i]
    why = sorted(((int(r[h.index(c)]), c[6:]) for c in stalls), reverse=True)[:2]
    print(f"{i:5d} {r[isamp]:>6} ex={r[iex]:>8} {why} {r[isrc][:80]}")
