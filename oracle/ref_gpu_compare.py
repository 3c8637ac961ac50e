"""TEST / MEASUREMENT INFRASTRUCTURE (oracle/), SURVEY 8f-3: the reference's own CUDA backends (wavefront, megakernel;
compiled unmodified for sm_100 by `make -C oracle refgpu`) and B200Renderer, behind the same `Renderer` interface and
the viewer's call pattern, on the same GPU. Prints one table; raw JSON lines go to the file given as argv[1].
The reference backends are fixed at depth 20, 1 spp per ProgressiveRender; B200Renderer runs at depth 20 with 1, 16
and 64 spp per call. Wall-clock per call (chrono around the loop, device synchronised on both sides)."""
import json
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DRV = os.path.join(HERE, "_ref", "ref_gpu_driver")
PRESETS = [(3, "CORNELL"), (1, "LIGHT_TEST"), (2, "MATERIAL_TEST"), (0, "DEFAULT"), (4, "RANDOM_BALLS_SMALL"),
           (5, "RANDOM_BALLS_MEDIUM"), (6, "RANDOM_BALLS_LARGE")]
W, H = 1920, 1080


def run(backend, preset, frames, spp, display=0):
    out = subprocess.run([DRV, backend, str(preset), str(W), str(H), "3", str(frames), str(spp), str(display)],
                         capture_output=True, text=True, timeout=600)
    line = [l for l in out.stdout.splitlines() if l.startswith("{")]
    if out.returncode != 0 or not line:
        return {"backend": backend, "preset": preset, "error": (out.stderr or out.stdout)[-300:]}
    return json.loads(line[-1])


def main():
    rows = []
    print("%-20s %5s | %-28s | %-28s | %-30s | %s" % ("preset (1080p, d20)", "prims", "ref wavefront ms  (spp/s)", "ref megakernel ms  (spp/s)",
                                                   "B200Renderer 1 spp/call", "B200Renderer 16 / 64 spp per call: spp/s"))
    for pid, name in PRESETS:
        big = pid >= 5
        r = {b: run(b, pid, 10 if big else 50, 1) for b in ("wavefront", "megakernel")}
        r["b1"] = run("b200", pid, 50, 1)
        r["b16"] = run("b200", pid, 20, 16)
        r["b64"] = run("b200", pid, 8, 64)
        r["b1d"] = run("b200", pid, 50, 1, display=1)
        r["wd"] = run("wavefront", pid, 10 if big else 50, 1, display=1)
        rows.extend(r.values())
        f = lambda x: "%8.3f (%8.1f)" % (x["ms_per_call"], x["spp_per_s"]) if "ms_per_call" in x else "error"
        best_ref = max(r[b].get("spp_per_s", 0.0) for b in ("wavefront", "megakernel"))
        print("%-20s %5s | %-28s | %-28s | %-22s x%-6.1f | %9.1f / %9.1f   mean L: ref %.4f / %.4f ours %.4f" % (
            name, r["b1"].get("prims", "?"), f(r["wavefront"]), f(r["megakernel"]), f(r["b1"]),
            r["b1"].get("spp_per_s", 0.0) / max(best_ref, 1e-9), r["b16"].get("spp_per_s", 0.0), r["b64"].get("spp_per_s", 0.0),
            r["wavefront"].get("mean_radiance", 0), r["megakernel"].get("mean_radiance", 0), r["b64"].get("mean_radiance", 0)))
        print("%-20s       | with film.UpdateDisplayGPU() per frame (the viewer's loop): ref wavefront %s, B200Renderer %s" % (
            "", f(r["wd"]), f(r["b1d"])))
    if len(sys.argv) > 1:
        with open(sys.argv[1], "w") as fh:
            for x in rows:
                fh.write(json.dumps(x) + "\n")


if __name__ == "__main__":
    main()
