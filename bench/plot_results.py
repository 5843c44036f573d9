"""Draws the round's headline results as PNG charts with PIL (matplotlib is not in this image): the counterpart of the
reference's experiment plots (/root/reference/pic/*.png, Readme.md:283-294).
  profiles/scaling_r1.png           images/s vs GPUs: this framework, torch DDP and the reference's DataParallel on the same boxes
  profiles/allreduce_sweep_r1.png   bus bandwidth vs bucket size at 8 GPUs: NCCL, one-shot, two-shot
Usage: python bench/plot_results.py   (reads the JSON files under profiles/)"""
import json
import os

from PIL import Image, ImageDraw, ImageFont

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROF = os.path.join(ROOT, "profiles")
COLORS = [(31, 119, 180), (255, 127, 14), (44, 160, 44), (214, 39, 40), (148, 103, 189)]


def _font(size):
    for p in ("/usr/share/fonts/truetype/dejavu/DejaVuSans.ttf", "/usr/share/fonts/dejavu/DejaVuSans.ttf"):
        if os.path.exists(p):
            return ImageFont.truetype(p, size)
    return ImageFont.load_default()


def grouped_bars(path, title, groups, series, values, ylabel, fmt="%.0f"):
    """values[s][g]: height of series s in group g (None = missing)."""
    W, H, L, R, T, B = 1100, 560, 90, 30, 60, 80
    img = Image.new("RGB", (W, H), "white")
    d = ImageDraw.Draw(img)
    f, fs = _font(16), _font(12)
    d.text((L, 18), title, fill="black", font=f)
    vmax = max(v for row in values for v in row if v is not None) * 1.12
    d.line([(L, T), (L, H - B), (W - R, H - B)], fill="black", width=2)
    for i in range(6):
        y = H - B - (H - B - T) * i / 5
        d.line([(L - 4, y), (W - R, y)], fill=(225, 225, 225))
        d.text((8, y - 7), fmt % (vmax * i / 5), fill="black", font=fs)
    d.text((8, T - 24), ylabel, fill="black", font=fs)
    gw = (W - L - R) / len(groups)
    bw = gw * 0.8 / len(series)
    for g, gname in enumerate(groups):
        for s in range(len(series)):
            v = values[s][g]
            if v is None:
                continue
            x0 = L + g * gw + gw * 0.1 + s * bw
            y0 = H - B - (H - B - T) * v / vmax
            d.rectangle([x0, y0, x0 + bw - 2, H - B], fill=COLORS[s % len(COLORS)])
            d.text((x0, y0 - 14), fmt % v, fill="black", font=fs)
        d.text((L + g * gw + gw * 0.35, H - B + 8), gname, fill="black", font=fs)
    for s, name in enumerate(series):
        d.rectangle([L + 10 + s * 250, H - 34, L + 24 + s * 250, H - 20], fill=COLORS[s % len(COLORS)])
        d.text((L + 30 + s * 250, H - 36), name, fill="black", font=fs)
    img.save(path)
    return path


def load(name):
    with open(os.path.join(PROF, name)) as fh:
        txt = fh.read().strip()
    return json.loads(txt.splitlines()[-1]) if not txt.startswith("[") else json.loads(txt)


def main():
    ours = {1: load("bench_1gpu_latest.json")["value"], 2: load("bench_2gpu_latest.json")["value"],
            4: load("bench_4gpu.json")["value"], 8: load("bench_8gpu.json")["value"]}
    ref8 = load("bench_ref_8gpu.json")
    # reference arm measured at 1, 2 and 8 GPUs (profiles/README.md): value = the better DataParallel variant
    ref_dp = {1: 10704.0, 2: 6007.0, 4: None, 8: ref8.get("value")}
    ref_ddp = {1: None, 2: 17850.0, 4: None, 8: ref8.get("torch_ddp_value")}
    gpus = [1, 2, 4, 8]
    grouped_bars(os.path.join(PROF, "scaling_r1.png"), "MobileNetV2, batch 128 per GPU, B200: images/s (device-timed, whole job)",
                 ["%d GPU%s" % (n, "" if n == 1 else "s") for n in gpus],
                 ["this framework (DDP, bf16)", "reference model, torch DDP (fp32)", "reference nn.DataParallel (fp32)"],
                 [[ours[n] for n in gpus], [ref_ddp[n] for n in gpus], [ref_dp[n] for n in gpus]], "images/s")
    rows = load("allreduce_sweep_8gpu.json")
    sizes = [r["MB"] for r in rows]
    grouped_bars(os.path.join(PROF, "allreduce_sweep_r1.png"), "fp32 gradient-bucket all-reduce on 8 x B200: bus bandwidth (GB/s)",
                 ["%d MB" % m for m in sizes], ["NCCL", "one-shot (peer loads)", "two-shot (peer loads/stores)"],
                 [[r["nccl_busGBs"] for r in rows], [r.get("oneshot_busGBs") for r in rows], [r["twoshot_busGBs"] for r in rows]], "GB/s")
    print("wrote profiles/scaling_r1.png, profiles/allreduce_sweep_r1.png")


if __name__ == "__main__":
    main()
