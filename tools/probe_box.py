"""One-off probes of the GPU box (run under gpurun; writes gpurun_out/probe_box.json):
  * codec SDKs: does libnvcuvid / libnvidia-encode / libnvjpeg load, and what does cuvidGetDecoderCaps say
    (VERDICT r01 item 10 -- a codec probe; no graphics stack involved);
  * host topology: CPUs, NUMA nodes, GPU <-> CPU affinity (for the pinned-H2D ceiling).
"""
import ctypes as C
import json
import os
import subprocess
import sys

out = {}


def sh(cmd):
    try:
        return subprocess.run(cmd, shell=True, capture_output=True, text=True, timeout=60).stdout.strip()
    except Exception as e:  # noqa: BLE001
        return f"ERR {e}"


class CUVIDDECODECAPS(C.Structure):
    _fields_ = [("eCodecType", C.c_int), ("eChromaFormat", C.c_int), ("nBitDepthMinus8", C.c_uint),
                ("reserved1", C.c_uint * 3), ("bIsSupported", C.c_ubyte), ("nNumNVDECs", C.c_ubyte),
                ("nOutputFormatMask", C.c_ushort), ("nMaxWidth", C.c_uint), ("nMaxHeight", C.c_uint),
                ("nMaxMBCount", C.c_uint), ("nMinWidth", C.c_ushort), ("nMinHeight", C.c_ushort),
                ("bIsHistogramSupported", C.c_ubyte), ("nCounterBitDepth", C.c_ubyte),
                ("nMaxHistogramBins", C.c_ushort), ("reserved3", C.c_uint * 10)]


def probe_codecs():
    res = {}
    for name in ("libnvcuvid.so.1", "libnvcuvid.so", "libnvidia-encode.so.1", "libnvjpeg.so.12"):
        try:
            C.CDLL(name)
            res[name] = "loads"
        except OSError as e:
            res[name] = f"dlopen failed: {e}"
    if res.get("libnvcuvid.so.1") == "loads":
        try:
            import torch
            torch.cuda.init()
            torch.zeros(1, device="cuda")
            lib = C.CDLL("libnvcuvid.so.1")
            caps = {}
            for cname, cid in (("MPEG4", 2), ("H264", 4), ("JPEG", 5), ("HEVC", 8), ("VP9", 10), ("AV1", 11)):
                c = CUVIDDECODECAPS()
                c.eCodecType, c.eChromaFormat, c.nBitDepthMinus8 = cid, 1, 0
                rc = lib.cuvidGetDecoderCaps(C.byref(c))
                caps[cname] = {"rc": rc, "supported": int(c.bIsSupported), "nvdecs": int(c.nNumNVDECs),
                               "max": [int(c.nMaxWidth), int(c.nMaxHeight)], "min": [int(c.nMinWidth), int(c.nMinHeight)]}
            res["cuvidGetDecoderCaps"] = caps
        except Exception as e:  # noqa: BLE001
            res["cuvidGetDecoderCaps"] = f"failed: {e}"
    return res


out["codecs"] = probe_codecs()
out["nproc"] = os.cpu_count()
out["affinity"] = sorted(os.sched_getaffinity(0))
out["numa_nodes_online"] = sh("cat /sys/devices/system/node/online")
out["lscpu"] = sh("lscpu | grep -E 'Model name|Socket|NUMA|^CPU\\(s\\)|Thread'")
out["topo"] = sh("nvidia-smi topo -m")
out["gpus"] = sh("nvidia-smi --query-gpu=index,name,pci.bus_id,pcie.link.gen.current,pcie.link.width.current --format=csv")
out["mem"] = sh("grep -E 'MemTotal|MemAvailable' /proc/meminfo")
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/probe_box.json", "w"), indent=1)
print(json.dumps(out, indent=1))
