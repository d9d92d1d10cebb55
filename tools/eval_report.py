#!/usr/bin/env python
"""SSIM report over two evaluation folders — the reference's eval.py (eval.py:1-67) for the headless build.

An evaluation run (RenderingManager::StartEvaluation, host/volrenderbase.cpp; reference renderingmanager.cpp:261-317,819-867) leaves
<dir>/eval.csv and <dir>/img/0000.png, 0001.png, ... (one per parameter-space sample).  eval.py compares the images of two such
folders pairwise with ImageMagick's `compare -metric SSIM` and writes an Image / SSIM table plus mean / min / max.  ImageMagick is
absent here: the metric is cpp_volume_rendering_b200.metrics.ssim (a restatement of ImageMagick's defaults).

  python tools/eval_report.py <eval dir A or its img/> <eval dir B or its img/> [--out ssim_comparison_results.csv] [--count 200]

Rows follow eval.py: one per index 0000 .. count-1 (default: as many as folder A holds), "Missing" where a file is absent, "Error"
where it cannot be read or the sizes differ.  The table is written as CSV (and as .xlsx when openpyxl is importable, as eval.py does).
"""
import argparse
import csv
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def img_dir(path):
    sub = os.path.join(path, "img")
    return sub if os.path.isdir(sub) else path


def read_png(path):
    from PIL import Image
    return np.asarray(Image.open(path).convert("RGB"))


def compare_folders(dir1, dir2, count=None):
    """Returns (rows [{'Image', 'SSIM'}], valid ssim values)."""
    from cpp_volume_rendering_b200 import metrics
    dir1, dir2 = img_dir(dir1), img_dir(dir2)
    if count is None:
        count = len([f for f in os.listdir(dir1) if f.endswith(".png")])
    rows, valid = [], []
    for i in range(count):
        name = f"{i:04d}.png"
        p1, p2 = os.path.join(dir1, name), os.path.join(dir2, name)
        if not (os.path.exists(p1) and os.path.exists(p2)):
            rows.append({"Image": name, "SSIM": "Missing"})
            continue
        try:
            a, b = read_png(p1), read_png(p2)
            if a.shape != b.shape:
                raise ValueError(f"sizes differ: {a.shape} vs {b.shape}")
            s = metrics.ssim(a, b)
            rows.append({"Image": name, "SSIM": s})
            valid.append(s)
        except Exception as e:  # noqa: BLE001  (eval.py records the failure and goes on)
            print(f"error on {name}: {e}", file=sys.stderr)
            rows.append({"Image": name, "SSIM": "Error"})
    return rows, valid


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("dir1")
    ap.add_argument("dir2")
    ap.add_argument("--out", default="ssim_comparison_results.csv")
    ap.add_argument("--count", type=int, default=None)
    args = ap.parse_args()
    rows, valid = compare_folders(args.dir1, args.dir2, args.count)
    with open(args.out, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=["Image", "SSIM"])
        w.writeheader()
        w.writerows(rows)
    try:
        import pandas as pd
        import openpyxl  # noqa: F401
        pd.DataFrame(rows).to_excel(os.path.splitext(args.out)[0] + ".xlsx", index=False)
    except ImportError:
        pass
    print(f"results saved to {args.out}")
    if valid:
        print(f"mean SSIM: {np.mean(valid):.4f}\nmin SSIM: {np.min(valid):.4f}\nmax SSIM: {np.max(valid):.4f}")


if __name__ == "__main__":
    main()
