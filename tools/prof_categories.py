"""Group the kernel rows of a tools/profile_step.py table: python tools/prof_categories.py profile.txt"""
import re
import sys


def parse(path):
    rows = []
    for line in open(path):
        parts = line.split()
        if len(parts) < 8 or not parts[-1].isdigit():
            continue
        name = line[:92].strip()
        if not (name.startswith("void") or "nvjet" in name or "cutlass" in name or "affb200::" in name or "Memcpy" in name
                or "Memset" in name or "cudnn" in name):
            continue
        t = parts[-5]   # Self CUDA
        if t.endswith("ms"):
            us = float(t[:-2]) * 1000
        elif t.endswith("us"):
            us = float(t[:-2])
        else:
            continue
        rows.append((us, int(parts[-1]), name))
    return rows


def cat(n):
    if "fused_attn" in n or "scatter_wsum" in n or "table_grad" in n:
        return "attention"
    if "nvjet" in n or "cutlass" in n or "splitK" in n or "gemm" in n.lower() or "cudnn" in n or "nhwc" in n:
        return "gemm/conv"
    if "affb200" in n:
        return "affb200 other"
    return "torch other"


def main():
    rows = parse(sys.argv[1])
    cats = {}
    for us, c, n in rows:
        a = cats.setdefault(cat(n), [0.0, 0])
        a[0] += us
        a[1] += c
    tot = sum(v[0] for v in cats.values())
    for k, v in sorted(cats.items(), key=lambda kv: -kv[1][0]):
        print("%-14s %7.2f ms %5.1f%% %5d launches" % (k, v[0] / 1000, 100 * v[0] / tot, v[1]))
    for k in cats:
        print("---", k)
        for us, c, n in sorted([r for r in rows if cat(r[2]) == k], reverse=True)[:int(sys.argv[2]) if len(sys.argv) > 2 else 14]:
            print("%8.0f us %4d  %s" % (us, c, n))


if __name__ == "__main__":
    main()
