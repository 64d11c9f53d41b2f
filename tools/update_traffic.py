"""profiles/traffic.json from an `ncu --set full` capture of the dominant kernel, tied to the kernel sources it was taken on
(bench.py prints `roofline.traffic` only when the hash matches the sources it runs):
  python tools/update_traffic.py gpurun_out/<capture>.ncu-rep <config id> "<what the capture is>" """
import csv
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402

rep, cfg, what = sys.argv[1], sys.argv[2], sys.argv[3]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]


def col(name):
    i = hdr.index(name)
    scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}.get(units[i], 1.0)
    return [float(r[i].replace(",", "")) * scale for r in rows[2:]]


rd, wr = col("dram__bytes_read.sum"), col("dram__bytes_write.sum")
fp64 = col("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active")
dur = col("gpu__time_duration.sum")
n = len(rd)
full = [i for i in range(n) if dur[i] > 0.5 * max(dur)]  # the near-empty last colour is not a representative launch
prof = ROOT / "profiles" / "traffic.json"
tj = json.loads(prof.read_text()) if prof.exists() else {}
tj[str(cfg)] = (sum(rd) + sum(wr)) / n
tj[f"fp64_pipe_pct_{cfg}"] = sum(fp64[i] for i in full) / len(full)
tj[f"_source_{cfg}"] = (f"{what}: (dram__bytes_read.sum + dram__bytes_write.sum) summed over the {n} launches captured / {n}; "
                        f"FP64 pipe averaged over the {len(full)} full launches; launches under ncu are isolated and cache-cold "
                        "(the positions are re-read from DRAM by every launch, which they are not back to back)")
tj[f"_csrc_sha256_{cfg}"] = bench.kernel_source_hash(int(cfg))
tj.pop("_csrc_sha256", None)
tj.pop("_source", None)
tj.pop("_fp64_source", None)
prof.write_text(json.dumps(tj, indent=1) + "\n")
print(json.dumps(tj, indent=1))
