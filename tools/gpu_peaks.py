"""Roofline denominators measured on the box (run under gpurun): L2 and HBM read bandwidth, SM count, issue / FP32 peaks."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "vulkan-pathtracer_b200"))
import b200pt

def main():
    out = {}
    for label, nbytes, iters in (("l2_16MiB", 16 << 20, 200), ("l2_32MiB", 32 << 20, 100), ("l2_64MiB", 64 << 20, 50),
                                 ("l2_96MiB", 96 << 20, 40), ("hbm_2GiB", 2 << 30, 3)):
        g, sms = b200pt.read_bandwidth(nbytes, iters)
        out[label + "_read_gbs"] = round(g, 1); out["num_sms"] = sms
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    mhz = peaks.get("sm_max_mhz", 1965.0)
    out["sm_max_mhz"] = mhz
    out["issue_peak_gwarpinst_s"] = round(out["num_sms"] * 4 * mhz * 1e-3, 1)
    out["fp32_peak_tflops"] = round(out["num_sms"] * 128 * 2 * mhz * 1e-6, 2)
    print(json.dumps(out))

if __name__ == "__main__":
    main()
