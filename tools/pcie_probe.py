"""Host<->device copy ceiling of the box: H2D alone, D2H alone, both at once (pinned memory, two streams).
The e2e number of bench.py is bounded by the duplex figure: one 4K 10-bit picture is 24.9 MB each way."""
import json
import sys

import torch


def main():
    mb = int(sys.argv[1]) if len(sys.argv) > 1 else 25
    n = 40
    dev = torch.device("cuda:0")
    hin = [torch.empty(mb << 20, dtype=torch.uint8).pin_memory() for _ in range(4)]
    hout = [torch.empty(mb << 20, dtype=torch.uint8).pin_memory() for _ in range(4)]
    d = [torch.empty(mb << 20, dtype=torch.uint8, device=dev) for _ in range(4)]
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def run(h2d, d2h):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s1)
        f0.record(s2)
        for i in range(n):
            if h2d:
                with torch.cuda.stream(s1):
                    d[i % 2].copy_(hin[i % 4], non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    hout[i % 4].copy_(d[2 + i % 2], non_blocking=True)
        e1.record(s1)
        f1.record(s2)
        torch.cuda.synchronize()
        ms = max(e0.elapsed_time(e1), f0.elapsed_time(f1))
        return n * (mb << 20) / (ms * 1e-3) / 1e9

    for _ in range(2):
        run(True, True)
    out = {"mb": mb, "h2d_GBps": round(run(True, False), 1), "d2h_GBps": round(run(False, True), 1),
           "duplex_each_GBps": round(run(True, True), 1)}
    print(json.dumps(out))


if __name__ == "__main__":
    main()
