"""profiles/r2_dram_traffic.json from an ncu summary (tools/ncu_summary.py --csv) of one 512^3 AirMix time step:
measured DRAM bytes per launch and per cell of each staged kernel (the `roofline.traffic` of bench.py).
    python tools/make_dram_traffic.py profiles/<tag>_staged_kernels_ncu_full_summary.csv [cells_per_direction]"""
import csv, json, os, sys

KEYS = [("euler_z_march", "euler_z_march_kernel"), ("euler_xy", "euler_xy_bulk_kernel"), ("diffusion_z", "diffusion_async_kernel<2"),
        ("diffusion_xy", "diffusion_async_kernel<1"), ("rk_stage", "rk_stage_rows_kernel")]


def main():
    src = sys.argv[1]
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 512
    rows = list(csv.reader(open(src)))
    hdr = rows[0]
    ik, ir, iw, it = (hdr.index(k) for k in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
    unit = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}
    ur, uw = unit[rows[1][ir]], unit[rows[1][iw]]
    out = {}
    for label, pat in KEYS:
        for r in rows[2:]:
            if pat in r[ik]:
                b = float(r[ir])*ur + float(r[iw])*uw
                out[label] = {"dram_bytes_per_cell": b/n**3, f"dram_bytes_per_launch_{n}": b, "duration_ms_under_ncu": float(r[it]),
                              "kernel": r[ik][:80]}
                break
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", "r2_dram_traffic.json")
    doc = {"AirMix": out, "source": f"{src} (ncu --set full --clock-control none, one RK3 time step of the air{n} box, "
                                      "tools/exp_timestep_once.py, one launch each; tools/measure_round3.sh)"}
    json.dump(doc, open(dst, "w"), indent=1)
    print(json.dumps({k: round(v["dram_bytes_per_cell"], 1) for k, v in out.items()}))


if __name__ == "__main__":
    main()
