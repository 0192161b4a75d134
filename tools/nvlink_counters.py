"""Why does an SM push top out near 713 GB/s when the copy engine writes 777 GB/s over the same NVLink port?
ncu's NVLink byte counters read 0 on this pool, so this tool brackets ~1 s loops of each way of moving the same bytes
GPU 0 -> GPU 1 with the driver's per-link counters (`nvidia-smi nvlink -gt d` = payload KiB, `-gt r` = raw KiB incl.
protocol overhead, where the driver offers it) and prints payload rate, raw rate and raw/payload per mode:

    push (TMA kernel)  |  push (register kernel)  |  pull (TMA kernel, run by GPU 1)  |  copy engine (cudaMemcpyPeerAsync)

ONE process, 2 GPUs.    python tools/nvlink_counters.py [--seconds 1.0] [--size 1073741824]
"""
import argparse
import json
import re
import subprocess
import sys
import time

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import torch

import dissem_b200 as d


def counters(kind: str, gpu: int):
    """{(link, 'tx'|'rx'): KiB} from `nvidia-smi nvlink -gt <kind> -i <gpu>`; raw text too."""
    try:
        out = subprocess.run(["nvidia-smi", "nvlink", "-gt", kind, "-i", str(gpu)], capture_output=True, text=True, timeout=20).stdout
    except Exception as e:                                   # noqa: BLE001
        return {}, f"failed: {e}"
    vals = {}
    for m in re.finditer(r"Link\s+(\d+):\s*(?:Raw|Data)\s+(Tx|Rx):\s*(\d+)\s*KiB", out):
        vals[(int(m.group(1)), m.group(2).lower())] = int(m.group(3))
    return vals, out


def delta(a, b, direction):
    return sum(b[k] - a[k] for k in b if k in a and k[1] == direction) * 1024


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=1.0)
    ap.add_argument("--size", type=int, default=1 << 30)
    ap.add_argument("--layers", type=int, default=4)
    args = ap.parse_args()
    if torch.cuda.device_count() < 2:
        print(json.dumps({"error": "needs 2 GPUs"}))
        return
    S, L = args.size, args.layers
    first = True
    for mode in ("push_tma", "push_register", "pull_tma", "copy_engine"):
        eng = d.Engine([0, 1])
        try:
            eng.set_option("variant", 0 if mode == "push_register" else 1)
            for l in range(L):
                eng.alloc(0, l, S)
                eng.fill(0, l, d.FILL_PATTERN, 0x5EED)
                eng.alloc(1, l, S)
            flags = d.XFER_PULL if mode == "pull_tma" else 0
            s = eng.compile([d.Xfer(0, 1, l, flags, 0, S, S) for l in range(L)])
            eng.wait(eng.launch(s))
            ce_src = ce_dst = None
            if mode == "copy_engine":
                # torch tensors of the same size on the two GPUs; .copy_ across devices is cudaMemcpyPeerAsync
                ce_src = [torch.empty(S, dtype=torch.uint8, device="cuda:0") for _ in range(L)]
                ce_dst = [torch.empty(S, dtype=torch.uint8, device="cuda:1") for _ in range(L)]
                for a, b in zip(ce_src, ce_dst):
                    b.copy_(a, non_blocking=True)
                torch.cuda.synchronize(0)
                torch.cuda.synchronize(1)
            d0, raw_text_d = counters("d", 0)
            r0, raw_text_r = counters("r", 0)
            d0_1, _ = counters("d", 1)
            t0 = time.perf_counter()
            moved = 0
            dev_ms = 0.0
            while time.perf_counter() - t0 < args.seconds:
                if mode == "copy_engine":
                    for a, b in zip(ce_src, ce_dst):
                        b.copy_(a, non_blocking=True)
                    torch.cuda.synchronize(0)
                    torch.cuda.synchronize(1)
                else:
                    tickets = [eng.launch(s) for _ in range(4)]
                    dev_ms += sum(eng.wait_detail(t)[0] for t in tickets)
                moved += L * S * (1 if mode == "copy_engine" else 4)
            wall = time.perf_counter() - t0
            d1, _ = counters("d", 0)
            r1, _ = counters("r", 0)
            d1_1, _ = counters("d", 1)
            row = {"mode": mode, "bytes_moved": moved, "wall_s": wall, "GBps_wall": moved / wall / 1e9,
                   "GBps_device": (moved / (dev_ms / 1e3) / 1e9) if dev_ms else None,
                   "gpu0_data_tx_bytes": delta(d0, d1, "tx"), "gpu0_data_rx_bytes": delta(d0, d1, "rx"),
                   "gpu1_data_tx_bytes": delta(d0_1, d1_1, "tx"), "gpu1_data_rx_bytes": delta(d0_1, d1_1, "rx"),
                   "gpu0_raw_tx_bytes": delta(r0, r1, "tx") if r0 else None, "gpu0_raw_rx_bytes": delta(r0, r1, "rx") if r0 else None,
                   "links_parsed": len({k[0] for k in d0})}
            if row["gpu0_data_tx_bytes"]:
                row["tx_payload_over_moved"] = row["gpu0_data_tx_bytes"] / moved
                row["rx_over_tx_on_gpu0"] = row["gpu0_data_rx_bytes"] / row["gpu0_data_tx_bytes"]
            if first:
                row["counter_text_sample_d"] = raw_text_d[:600]
                row["counter_text_sample_r"] = raw_text_r[:600]
                first = False
            print(json.dumps(row), flush=True)
            eng.sched_free(s)
        finally:
            eng.close()


if __name__ == "__main__":
    main()
