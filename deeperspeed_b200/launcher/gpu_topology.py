"""NVLink pair detection from `nvidia-smi topo -m` (DeeperSpeed addition, reference launcher/gpu_topology.py:7-79).
On NVSwitch systems (B200 HGX) every pair is NV18 so any order is optimal; on PCIe / partial-NVLink boxes the local
ranks are re-ordered so that adjacent pipeline stages sit on NVLink-connected GPUs."""
import subprocess


def nvidia_smi_topo():
    try:
        return subprocess.check_output(["nvidia-smi", "topo", "-m"], text=True)
    except Exception:
        return ""


def parse_topology(text):
    """Returns {gpu_i: {gpu_j: link}} from the matrix printed by nvidia-smi topo -m."""
    lines = [l for l in text.splitlines() if l.strip()]
    header = None
    matrix = {}
    for l in lines:
        cols = l.split()
        if header is None:
            if cols and cols[0].startswith("GPU0"):
                header = [c for c in cols if c.startswith("GPU")]
            continue
        if not cols[0].startswith("GPU"):
            continue
        i = int(cols[0][3:])
        matrix[i] = {}
        for j, link in enumerate(cols[1:1 + len(header)]):
            matrix[i][j] = link
    return matrix


def detect_nvlink_pairs(text=None):
    """Greedy pairing of GPUs connected by NVLink (NV# entries); returns a flat device order, e.g. [0, 3, 1, 2]."""
    m = parse_topology(text if text is not None else nvidia_smi_topo())
    if not m:
        return []
    unpaired = sorted(m.keys())
    order = []
    while unpaired:
        i = unpaired.pop(0)
        order.append(i)
        best, best_w = None, -1
        for j in unpaired:
            link = m[i].get(j, "")
            w = int(link[2:]) if link.startswith("NV") and link[2:].isdigit() else -1
            if w > best_w:
                best, best_w = j, w
        if best is not None and best_w > 0:
            unpaired.remove(best)
            order.append(best)
    return order
