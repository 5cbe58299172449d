"""codes_to_wavs post-processing alone (for ncu launch lists): python scripts/bench_postproc.py [batch] [frames] [chunk,chunk,...]."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

if __name__ == "__main__":
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    frames = int(sys.argv[2]) if len(sys.argv) > 2 else 2580
    peaks = bench.peaks()
    hbm = peaks[0] if isinstance(peaks, tuple) else None
    chunks = [int(c) for c in sys.argv[3].split(",")] if len(sys.argv) > 3 else [None]
    for chunk in chunks:
        r = bench.postproc_section(torch.device("cuda:0"), hbm, batch, frames, chunk=chunk)
        r["chunk"] = chunk
        print(json.dumps(r), flush=True)
