"""Regenerate profiles/sass/: per-kernel SASS listings and counts of the instructions that carry
the fused compute+communication protocol (system-scope flag traffic, reductions, fences).

    python tools/sass_report.py            # needs cuobjdump (CUDA toolkit), no GPU
"""
import collections
import re
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
LIB = ROOT / "parameter_server_b200" / "lib" / "libpsb200_cuda.so"
OUT = ROOT / "profiles" / "sass"
KERNELS = {
    "k_darlin_passILb0ELb1": "k_darlin_pass_single.sass",
    "k_darlin_passILb0ELb0": "k_darlin_pass_multi.sass",
    "k_sgd_worker_step": "k_sgd_worker_step.sass",
    "k_sgd_server_all": "k_sgd_server_all.sass",
    "k_sgd_serverE": "k_sgd_server.sass",
    "k_kv_bulk_pull_apply": "k_kv_bulk_pull_apply.sass",
    "k_kv_bulk_gather_send32": "k_kv_bulk_gather_send32.sass",
    "k_kv_push_add": "k_kv_push_add.sass",
    "k_dense_marginsILi64": "k_dense_margins_n64.sass",
    "k_dkv_applyId": "k_dkv_apply_f64.sass",
}
INTEREST = re.compile(r"\b(ATOMG|ATOM|REDG|RED|MEMBAR|ERRBAR|CGAERRBAR|CCTL|LDG\.E[\w.]*STRONG[\w.]*|"
                      r"STG\.E[\w.]*STRONG[\w.]*|LDG\.E\.128[\w.]*|STG\.E\.128[\w.]*|DFMA|DADD|DMUL|MUFU[\w.]*|"
                      r"UTC\w*MMA|UTMALDG|UBLKCP|UTCBAR|LDTM|STTM|SYNCS|HMMA|SHFL\.\w+)[\w.]*")


def main() -> int:
    txt = subprocess.run(["cuobjdump", "-sass", str(LIB)], capture_output=True, text=True, check=True).stdout
    funcs = re.split(r"\n\s*Function : ", txt)[1:]
    OUT.mkdir(parents=True, exist_ok=True)
    lines = ["# SASS evidence (cuobjdump -sass libpsb200_cuda.so, sm_100a)",
             "Regenerate with `python tools/sass_report.py`.  Counts of the instructions that implement the",
             "fused compute+communication protocol: `*.STRONG.SYS` = system-scope acquire/release or volatile",
             "traffic on peer-visible memory (tagged landing entries, flags), `RED`/`ATOM` = reductions / atomics",
             "(margin scatter, counters, CAS inserts), `MEMBAR.*.SYS` + `ERRBAR` = fences before cross-GPU",
             "stores, `.STRONG.GPU` = grid-barrier traffic, `CCTL.IVALL` = L1 invalidation after an acquire,",
             "`DFMA`/`MUFU.RCP64H` = the fp64 exp / reciprocal of the logistic gradient.",
             "The sparse kernels carry no tensor-core instruction (SpMV-class work, see DESIGN.md); the dense",
             "feature-block tile does: `UTCHMMA` = tcgen05.mma, `UTMALDG` = TMA tensor load, `LDTM` = tcgen05.ld,",
             "`UTCBAR` = tcgen05.commit, `SYNCS` = mbarrier operations.", ""]
    for f in funcs:
        name = f.split("\n", 1)[0].strip()
        for key, fn in KERNELS.items():
            if key in name:
                body = f
                # keep address + instruction, drop the hex encodings (8x smaller listings)
                slim = [re.sub(r"\s*/\* 0x[0-9a-f]{16} \*/\s*$", "", ln) for ln in body.splitlines()]
                (OUT / fn).write_text("Function : " + "\n".join(ln for ln in slim if ln.strip()) + "\n")
                ins = re.findall(r"^\s+/\*[0-9a-f]+\*/\s+(.*?);", body, flags=re.M)
                cnt = collections.Counter()
                for i in ins:
                    m = INTEREST.search(i)
                    if m:
                        cnt[m.group(0)] += 1
                lines.append(f"## `{name[:96]}`")
                lines.append(f"{len(ins)} SASS instructions -> `{fn}`\n")
                lines += [f"- `{k}` x{v}" for k, v in sorted(cnt.items())]
                lines.append("")
    (OUT / "README.md").write_text("\n".join(lines))
    print("wrote", OUT)
    return 0


if __name__ == "__main__":
    sys.exit(main())
