"""Writes the per-kernel evidence under profiles/:
  * profiles/sass/<kernel>.sass      -- cuobjdump -sass of one instantiation per kernel family
  * profiles/sass_summary.md         -- per kernel: registers, tcgen05 / TMA / peer mnemonic counts
  * profiles/ncu_<name>.md           -- key metrics read from gpurun_out/*.ncu-rep (if present)
Runs on the CPU box (needs only the built .so and ncu for reading reports)."""
import csv
import io
import re
import subprocess
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
SO = ROOT / "simple_distributed_machine_learning_b200" / "_lib" / "libsdml_b200.so"
OUT = ROOT / "profiles"
(OUT / "sass").mkdir(parents=True, exist_ok=True)

# Preferred instantiation of templated kernels (the one the training path launches); every other
# __global__ function found in csrc/*.cu is listed with the first instantiation cuobjdump shows.
PREFERRED = {
    "gemm_bf16_2cta_kernel": r"gemm_bf16_2cta_kernelILi256ELi6ELb0ELb0",
    "gemm_bf16_2cta_comm_kernel": r"gemm_bf16_2cta_comm_kernelILi256ELi6ELb0ELb0",
    "gemm_bf16_kernel": r"gemm_bf16_kernelILi256ELi4ELb0ELb0",
    "gemm_mxfp8_2cta_kernel": r"gemm_mxfp8_2cta_kernelILi192ELb0",
    "softmax_ce_rows_kernel": r"softmax_ce_rows_kernelI13__nv_bfloat16Li4",
    "softmax_ce_kernel": r"softmax_ce_kernelI13__nv_bfloat16",
}
# extra listings of a second instantiation that matters on its own
EXTRA = {"gemm_bf16_2cta_kernel__wgrad": r"gemm_bf16_2cta_kernelILi256ELi6ELb1ELb1"}


def global_kernels():
    """Names of every __global__ function defined in csrc/*.cu."""
    names = []
    for src in sorted((ROOT / "csrc").glob("*.cu")):
        text = src.read_text()
        for m in re.finditer(r"__global__\s+void(?:\s+__cluster_dims__\([^)]*\))?(?:\s+__launch_bounds__\([^)]*\))?\s+(\w+)\s*\(", text):
            names.append((m.group(1), src.name))
    return names


MNEMONICS = ["UTCHMMA", "UTCQMMA", "UTCCP", "UTCBAR", "LDTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP",
             "SYNCS", "HMMA", "ATOM", "RED", "LDG", "STG", "MEMBAR", "ELECT"]


def main():
    sass = subprocess.run(["cuobjdump", "-sass", str(SO)], capture_output=True, text=True).stdout
    funcs = re.split(r"\n\s*Function : ", sass)
    by_name = {}
    for f in funcs[1:]:
        name = f.split("\n", 1)[0].strip()
        by_name[name] = f
    res = subprocess.run(["cuobjdump", "-res-usage", str(SO)], capture_output=True, text=True).stdout
    regs = {}
    cur = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            cur = m.group(1)
        m = re.search(r"REG:(\d+).*SHARED:(\d+)", line)
        if m and cur:
            regs[cur] = (int(m.group(1)), int(m.group(2)))
    lines = ["# SASS evidence (cuobjdump -sass of libsdml_b200.so, sm_100a)", "",
             "`tcgen05.mma` -> UTCHMMA / UTCQMMA (block-scaled fp8), `tcgen05.ld` -> LDTM, `tcgen05.cp` -> UTCCP, "
             "`tcgen05.commit` -> UTCBAR, TMA loads / stores / reduce-stores -> UTMALDG / UTMASTG / UTMAREDG, "
             "bulk copies -> UBLKCP, mbarrier ops -> SYNCS.  No HMMA (legacy mma.sync) anywhere.", "",
             "| kernel | registers | static smem | " + " | ".join(MNEMONICS) + " |",
             "|---|---|---|" + "---|" * len(MNEMONICS)]
    kernels = global_kernels()
    families = {k: PREFERRED.get(k, r"\d+" + k + r"(?:I|E|P|N|v|[a-z])") for k, _ in kernels}
    families.update(EXTRA)
    for old in (OUT / "sass").glob("*.sass"):
        old.unlink()
    for fam, pat in families.items():
        cands = [n for n in by_name if re.search(pat, n)]
        if not cands:
            lines.append(f"| {fam} | (not found) |")
            continue
        name = sorted(cands)[0]
        body = by_name[name]
        (OUT / "sass" / f"{fam}.sass").write_text("Function : " + body)
        counts = [len(re.findall(r"\b" + m + r"[.A-Z0-9_]*\b", body)) for m in MNEMONICS]
        r = regs.get(name, ("?", "?"))
        lines.append(f"| {fam} (`{name[:60]}`) | {r[0]} | {r[1]} | " + " | ".join(str(c) for c in counts) + " |")
    (OUT / "sass_summary.md").write_text("\n".join(lines) + "\n")
    print("\n".join(lines[-len(families):]))

    want = ["gpu__time_duration.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
            "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
            "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
            "l1tex__m_xbar2l1tex_read_bytes.sum", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "sm__cycles_elapsed.avg.per_second", "launch__grid_size", "launch__block_size",
            "launch__shared_mem_per_block_dynamic", "sm__inst_executed_pipe_tensor.sum",
            "smsp__cycles_active.avg", "l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum",
            "l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum", "lts__t_bytes.sum",
            "nvlrx__bytes.sum", "nvltx__bytes.sum", "sm__warps_active.avg.per_cycle_active"]
    for rep in sorted((ROOT / "gpurun_out").glob("*.ncu-rep")):
        raw = subprocess.run(["ncu", "-i", str(rep), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        if len(rows) < 3:
            continue
        hdr, units = rows[0], rows[1]
        out = [f"# ncu --set full: {rep.name}", ""]
        seen = {}
        for vals in rows[2:]:
            kname = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
            seen[kname] = seen.get(kname, 0) + 1
            if seen[kname] > 2:          # the first two launches of a kernel are enough
                continue
            out += [f"## {kname[:120]}", "", "| metric | value | unit |", "|---|---|---|"]
            for h, u, v in zip(hdr, units, vals):
                if h in want:
                    out.append(f"| {h} | {v} | {u} |")
            out.append("")
        (OUT / f"ncu_{rep.stem}.md").write_text("\n".join(out))
        print("wrote", OUT / f"ncu_{rep.stem}.md")


if __name__ == "__main__":
    main()
