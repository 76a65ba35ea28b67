#!/usr/bin/env python
"""overlap_probe.py — does running two half-batches on two streams (stages of different batches side by side: FAST is
ALU-bound, the blur waits on loads, the quadtree on barriers, the descriptor kernel on the L2 path) beat one batch whose
stages run one after the other?  Prints frames/s for one handle x 1024 frames and for H handles x 1024/H frames, each on
its own stream, free running, with a start offset between them.

  python tools/overlap_probe.py [--frames 1024] [--steps 10]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import orb_slam2_annotate_b200 as orbfe
    from orb_slam2_annotate_b200 import synth
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=1024)
    ap.add_argument("--steps", type=int, default=10)
    args = ap.parse_args()
    W, H, B = 640, 480, args.frames
    base, _ = synth.sequence(64, W, H, seed=1)
    frames = np.ascontiguousarray(np.tile(base, (B // 64 + 1, 1, 1))[:B])
    d_frames = torch.from_numpy(frames).cuda()
    out = {}
    for nh in (1, 2, 3, 4):
        per = B // nh
        exs = [orbfe.ORBextractor(1000, max_width=W, max_height=H, max_batch=per) for _ in range(nh)]
        streams = [torch.cuda.ExternalStream(e.stream) for e in exs]
        ptrs = [d_frames[i * per:(i + 1) * per].data_ptr() for i in range(nh)]
        for delay_us in ((0,) if nh == 1 else (0, 400)):
            def run(n):
                for s_ in range(n):
                    for i, e in enumerate(exs):
                        if s_ == 0 and delay_us and i:
                            with torch.cuda.stream(streams[i]):
                                torch.cuda._sleep(int(delay_us * i * 1900))
                        e.extract_batch_device(ptrs[i], per, W, H)
            run(3)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for s in streams:
                s.wait_event(e0)
            run(args.steps)
            done = [torch.cuda.Event() for _ in streams]
            for s, ev in zip(streams, done):
                ev.record(s)
                torch.cuda.current_stream().wait_event(ev)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.steps
            out[f"handles{nh}_delay{delay_us}"] = {"ms_per_1024": ms * 1024 / (per * nh), "frames_per_s": per * nh / ms * 1e3}
        del exs
    print(json.dumps(out))


if __name__ == "__main__":
    main()
