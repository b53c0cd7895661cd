"""dram__bytes_read.sum + dram__bytes_write.sum per launch of every kernel group, from an `ncu --page raw --csv` export of a
`--set full` capture of `bench.py` (512^3 x 180, 1 GPU) -> profiles/traffic_512.json (read by bench.py's `roofline.traffic`).
usage: python scripts/ncu_traffic.py <raw.csv> [out.json]"""
import csv, json, os, sys

GROUPS = [("fgp2_kernel<1, 0>", "fgp2_first"), ("fgp2_kernel<0, 0>", "fgp2_sweep"), ("fgp2_kernel<0, 2>", "fgp2_final"), ("fgp2_kernel<1, 2>", "fgp2_first_final"),
          ("fgp_tile_kernel<1", "fgp_first"), ("fgp_tile_kernel<0", "fgp_sweep"), ("fgp_tile_kernel<2", "fgp_final"),
          ("fpt_kernel", "fp"), ("fp_kernel", "fp"), ("bpt_kernel", "bp"), ("bp_kernel", "bp"), ("tv_value", "tv"), ("momentum", "momentum"),
          ("tvgd_gradient", "tvgd_grad"), ("tvgd_update", "tvgd_update")]


def unit_scale(u):
    return {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(u, 1.0)


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    ix = {k: i for i, k in enumerate(hdr)}
    out = {}
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        name = r[ix["Kernel Name"]]
        grp = next((g for key, g in GROUPS if key in name), None)
        if grp is None or grp in out:
            continue
        rd = float(r[ix["dram__bytes_read.sum"]]) * unit_scale(units[ix["dram__bytes_read.sum"]])
        wr = float(r[ix["dram__bytes_write.sum"]]) * unit_scale(units[ix["dram__bytes_write.sum"]])
        out[grp] = rd + wr
    dst = sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic_512.json")
    json.dump(out, open(dst, "w"), indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
