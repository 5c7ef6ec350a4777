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


# ---- the reference's function names (launcher/gpu_topology.py:7-79), built on the parser above ---------------------
def get_topology_str():
    return nvidia_smi_topo()


def is_nvlink(connection_type):
    """NV<n> entries of the topology matrix are NVLink connections."""
    return connection_type.startswith("NV") and connection_type[2:].isdigit()


def topology_from_string(string):
    """Topology matrix as a list of rows (row i = links of GPU i to GPU 0..n-1)."""
    m = parse_topology(string)
    n = len(m)
    rows = [[m[i].get(j, "X") for j in range(n)] for i in sorted(m)]
    assert all(len(r) == len(rows) for r in rows)
    return rows


def contains_nvlinks(topology):
    return any(is_nvlink(item) for row in topology for item in row)


def get_nvlink_pairs(topology):
    """Set of (i, j), i < j in first-seen order, of GPU pairs bridged by NVLink."""
    out = set()
    for i, row in enumerate(topology):
        for j, link in enumerate(row):
            if is_nvlink(link) and (j, i) not in out:
                out.add((i, j))
    return out


def get_cuda_visible_device_mapping(nvlink_pairs, local_gpu_ids=None):
    """CUDA_VISIBLE_DEVICES string that puts NVLink partners next to each other."""
    flat = [g for pair in sorted(nvlink_pairs) for g in pair]
    if local_gpu_ids is not None:
        flat = [g for g in flat if g in local_gpu_ids]
    seen = []
    for g in flat:
        if g not in seen:
            seen.append(g)
    return ",".join(map(str, seen))


def detect_nvlink_pairs_and_map_visible_devices(rank, local_gpu_ids):
    topology = topology_from_string(get_topology_str())
    if topology and contains_nvlinks(topology):
        return get_cuda_visible_device_mapping(get_nvlink_pairs(topology), local_gpu_ids)
    print(f"No NVLINK detected on rank {rank}")
    return None
