"""The hardware view of one kernel launch as JSON (bench.py's roofline.hardware): `ncu --set full` report -> the counters
SURVEY.md 8(d) asks to report beside the algorithmic roofline.
usage: ncu_hw_json.py <report.ncu-rep> "<what was captured>" > profiles/r2_mega_hw_4k64.json"""
import csv, json, subprocess, sys

rep, what = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, r = rows[0], rows[1], rows[2]
d = dict(zip(hdr, r))
u = dict(zip(hdr, units))


def f(name):
    try:
        return float(d[name].replace(",", ""))
    except Exception:
        return None


def scaled(name, unit_to):
    """ncu prints durations / byte counts in an adaptive unit: bring them to a fixed one."""
    v = f(name)
    if v is None:
        return None
    unit = u.get(name, "")
    mult = {"ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "s": 1.0, "second": 1.0, "nsecond": 1e-9,
            "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1.0)
    return v * mult / unit_to


dur_s = scaled("gpu__time_duration.sum", 1.0)
warp_inst = f("smsp__inst_executed.sum")
ratio = f("smsp__thread_inst_executed_per_inst_executed.ratio")
thread_inst = f("smsp__thread_inst_executed.sum") or (warp_inst * ratio if warp_inst and ratio else None)
o = {
    "source": f"ncu --set full --clock-control none, one launch: {what}",
    "kernel": d.get("Kernel Name"),
    "duration_ms_under_ncu": None if dur_s is None else dur_s * 1e3,
    "registers_per_thread": f("launch__registers_per_thread"),
    "smsp__issue_active_pct": f("smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "thread_inst_executed_per_inst_executed": f("smsp__thread_inst_executed_per_inst_executed.ratio"),
    "warp_inst_executed": warp_inst,
    "thread_inst_executed": thread_inst,
    "executed_Tlane_instr_per_s": None if not (thread_inst and dur_s) else thread_inst / dur_s / 1e12,
    "pipe_fma_pct": f("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
    "pipe_alu_pct": f("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    "pipe_xu_pct": f("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
    "pipe_lsu_pct": f("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
    "l1_hit_pct": f("l1tex__t_sector_hit_rate.pct"),
    "l2_hit_pct": f("lts__t_sector_hit_rate.pct"),
    "lts__t_bytes": scaled("lts__t_bytes.sum", 1.0),  # not in --set full on this ncu; the two L1<->L2 figures below stand in
    "l2_to_l1_read_bytes": scaled("l1tex__m_xbar2l1tex_read_bytes.sum", 1.0),
    "l1_to_l2_write_bytes": scaled("l1tex__m_l1tex2xbar_write_bytes.sum", 1.0),
    "dram_bytes_read": scaled("dram__bytes_read.sum", 1.0),
    "dram_bytes_write": scaled("dram__bytes_write.sum", 1.0),
    "local_loads": f("sass__inst_executed_local_loads"),
    "local_stores": f("sass__inst_executed_local_stores"),
    "warps_active_pct": f("sm__warps_active.avg.pct_of_peak_sustained_active"),
}
print(json.dumps(o, indent=1))
