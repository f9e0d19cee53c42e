#!/usr/bin/env python
"""What the box's PCIe link gives the host-pointer lookup(): H2D alone, D2H alone and both at
once, pinned and pageable, at the byte counts of one C2 call (78 MB in, 109 MB out).
Measurement aid for DESIGN.md section 6 (the e2e number is PCIe-bound); prints JSON lines.
"""
import json
import time

import torch


def bw(nbytes, ms):
    return round(nbytes / ms / 1e6, 2)


def main():
    torch.cuda.set_device(0)
    n_in, n_out = 78_405_492, 109_051_904
    hin = torch.empty(n_in, dtype=torch.uint8).pin_memory()
    hout = torch.empty(n_out, dtype=torch.uint8).pin_memory()
    pin = torch.empty(n_in, dtype=torch.uint8)
    pout = torch.empty(n_out, dtype=torch.uint8)
    din = torch.empty(n_in, dtype=torch.uint8, device="cuda")
    dout = torch.empty(n_out, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def timed(fn, reps=10):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / reps * 1e3

    def h2d(src=hin):
        with torch.cuda.stream(s1):
            din.copy_(src, non_blocking=True)

    def d2h(dst=hout):
        with torch.cuda.stream(s2):
            dst.copy_(dout, non_blocking=True)

    def both():
        h2d()
        d2h()

    def chunked(k):
        def f():
            ci, co = n_in // k, n_out // k
            for i in range(k):
                with torch.cuda.stream(s1):
                    din[i * ci:(i + 1) * ci].copy_(hin[i * ci:(i + 1) * ci], non_blocking=True)
                with torch.cuda.stream(s2):
                    hout[i * co:(i + 1) * co].copy_(dout[i * co:(i + 1) * co], non_blocking=True)
        return f

    ms = timed(h2d)
    print(json.dumps({"what": "h2d pinned alone", "ms": round(ms, 3), "GBps": bw(n_in, ms)}))
    ms = timed(d2h)
    print(json.dumps({"what": "d2h pinned alone", "ms": round(ms, 3), "GBps": bw(n_out, ms)}))
    ms = timed(both)
    print(json.dumps({"what": "h2d+d2h pinned concurrent", "ms": round(ms, 3), "GBps_sum": bw(n_in + n_out, ms),
                      "GBps_d2h": bw(n_out, ms)}))
    for k in (4, 26, 104):
        ms = timed(chunked(k))
        print(json.dumps({"what": f"concurrent, {k} chunks each way", "ms": round(ms, 3),
                          "GBps_sum": bw(n_in + n_out, ms)}))
    ms = timed(lambda: h2d(pin), reps=4)
    print(json.dumps({"what": "h2d pageable alone", "ms": round(ms, 3), "GBps": bw(n_in, ms)}))
    ms = timed(lambda: d2h(pout), reps=4)
    print(json.dumps({"what": "d2h pageable alone", "ms": round(ms, 3), "GBps": bw(n_out, ms)}))
    # host memcpy speed (staging cost if pageable buffers are copied through pinned ones)
    t0 = time.perf_counter()
    for _ in range(5):
        hin.copy_(pin)
    dt = (time.perf_counter() - t0) / 5 * 1e3
    print(json.dumps({"what": "host memcpy pageable->pinned (1 thread)", "ms": round(dt, 3), "GBps": bw(n_in, dt)}))


if __name__ == "__main__":
    main()
