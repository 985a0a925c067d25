"""Median bar chart.  matplotlib when it is installed, a Pillow drawing when it
is not (the GPU boxes have no matplotlib), a text table as the last resort."""
from __future__ import annotations

from typing import List


def _with_matplotlib(labels, values, legend, path) -> bool:
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except Exception:  # noqa: BLE001
        return False
    fig, ax = plt.subplots(figsize=(16, 6))
    bars = ax.bar(labels, values, color="skyblue")
    for bar, v in zip(bars, values):
        ax.text(bar.get_x() + bar.get_width() / 2, bar.get_height(), f"{v:.5f}", ha="center", va="bottom")
    ax.text(1.02, 0.95, legend, transform=ax.transAxes, fontsize=10, va="top",
            bbox=dict(facecolor="white", alpha=0.5))
    ax.set_xlabel("Device and Kernel Size")
    ax.set_ylabel("Median Execution Time (ms)")
    ax.set_title("Median Execution Time by Device and Kernel Size")
    fig.tight_layout()
    fig.savefig(path, dpi=300, bbox_inches="tight")
    plt.close(fig)
    return True


def _with_pillow(labels, values, legend, path) -> bool:
    try:
        from PIL import Image, ImageDraw
    except Exception:  # noqa: BLE001
        return False
    W, H, L, B, T = 1600, 600, 90, 110, 50
    plot_w = 1000
    img = Image.new("RGB", (W, H), "white")
    d = ImageDraw.Draw(img)
    vmax = max(values) if values and max(values) > 0 else 1.0
    d.text((L, 15), "Median Execution Time by Device and Kernel Size (ms)", fill="black")
    d.line([(L, T), (L, H - B), (L + plot_w, H - B)], fill="black")
    n = max(1, len(values))
    slot = plot_w / n
    for i, (lab, v) in enumerate(zip(labels, values)):
        x0 = L + i * slot + slot * 0.15
        x1 = L + (i + 1) * slot - slot * 0.15
        y = H - B - (H - B - T) * (v / vmax) * 0.92
        d.rectangle([x0, y, x1, H - B], fill=(135, 206, 235), outline="black")
        d.text((x0, y - 12), f"{v:.5f}", fill="black")
        d.text((x0, H - B + 6 + 12 * (i % 3)), str(lab)[:40], fill="black")
    d.multiline_text((L + plot_w + 30, T), legend[:3000], fill="black")
    img.save(path)
    return True


def median_bar_chart(labels: List[str], values: List[float], legend: str, path: str) -> str:
    if _with_matplotlib(labels, values, legend, path) or _with_pillow(labels, values, legend, path):
        return path
    txt = path.rsplit(".", 1)[0] + ".txt"
    with open(txt, "w") as f:
        for lab, v in zip(labels, values):
            f.write(f"{lab}\t{v:.5f}\n")
        f.write("\n" + legend)
    return txt
