"""Generates tests/golden/*.npz from the CPU oracle (the reference ships no golden vectors: SURVEY.md §4, parity unpinned).
Run from the repo root: python tests/golden/make_golden.py. The fixtures pin the oracle against regressions and are the
vectors the -m gpu tests check the CUDA path against on the GPU box (where /root/reference does not exist)."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from box2d_b200 import api, scenes  # noqa: E402

CASES = {
    "pyramid10": (lambda w: scenes.add_pyramid(w, 10), 10, {}),
    "mixed_drop": (lambda w: scenes.mixed_drop(w), 30, {}),
    "joint_grid8": (lambda w: scenes.joint_grid(w, 8), 20, {}),
    "tumbler300": (lambda w: scenes.tumbler(w, n=300, half=4.0), 20, {}),
    "joint_zoo": (lambda w: scenes.joint_zoo(w, copies=2), 40, {}),                       # every joint type (weld/prismatic use Atan2)
    "joint_zoo_exact": (lambda w: scenes.joint_zoo(w, copies=2, angles=False), 40, {}),   # the Atan2-free subset
    "wheel_rig": (lambda w: scenes.wheel_rig(w, copies=2), 30, {}),
    "bounce_mix": (lambda w: scenes.bounce_mix(w), 45, {}),                                # restitution != 0 (quirk Q1) and hit events
}


def record(world, build, steps):
    build(world)
    out = {}
    for s in range(steps):
        world.Step(1.0 / 60.0, 4)
        c = world.ReadContacts()
        ev = world.GetContactEvents()
        key = (np.minimum(c["shapeIdA"], c["shapeIdB"]).astype(np.uint64) << np.uint64(32)) | np.maximum(c["shapeIdA"], c["shapeIdB"]).astype(np.uint64)
        out[f"s{s}_pairs"] = np.sort(key)
        out[f"s{s}_contact"] = np.stack([c["contactId"], c["shapeIdA"], c["shapeIdB"], c["colorIndex"], c["pointCount"],
                                         c["points"]["id"][:, 0].astype(np.int32), c["points"]["id"][:, 1].astype(np.int32)], axis=1)
        out[f"s{s}_begin"] = np.stack([ev[0]["shapeIdA"], ev[0]["shapeIdB"], ev[0]["contactId"]], axis=1) if len(ev[0]) else np.zeros((0, 3), np.int32)
        out[f"s{s}_end"] = np.stack([ev[1]["shapeIdA"], ev[1]["shapeIdB"], ev[1]["contactId"]], axis=1) if len(ev[1]) else np.zeros((0, 3), np.int32)
        out[f"s{s}_local"] = np.asarray(c["localIndex"], dtype=np.int32)          # position in the colour's contactSims array (restitution groups, hit order)
        h = ev[2]
        out[f"s{s}_hit_ids"] = np.stack([h["shapeIdA"], h["shapeIdB"]], axis=1) if len(h) else np.zeros((0, 2), np.int32)
        out[f"s{s}_hit_data"] = (np.stack([h["point"]["x"], h["point"]["y"], h["normal"]["x"], h["normal"]["y"], h["approachSpeed"]], axis=1).astype(np.float32)
                                 if len(h) else np.zeros((0, 5), np.float32))
        if s == 0 or s == steps - 1:
            j = world.ReadJoints()
            if len(j):
                out[f"s{s}_joint_ids"] = np.stack([j["jointId"], j["colorIndex"], j["type"]], axis=1)
                out[f"s{s}_joint_impulses"] = np.stack([j["linearImpulse"]["x"], j["linearImpulse"]["y"], j["springImpulse"], j["motorImpulse"], j["lowerImpulse"],
                                                        j["upperImpulse"], j["linearImpulse2"]["x"], j["linearImpulse2"]["y"]], axis=1).astype(np.float32)
                out[f"s{s}_joint_events"] = np.asarray(world.GetJointEvents(), dtype=np.int32)
            b = world.ReadBodies()
            out[f"s{s}_bodies"] = np.stack([b["transform"]["p"]["x"], b["transform"]["p"]["y"], b["transform"]["q"]["c"], b["transform"]["q"]["s"],
                                            b["linearVelocity"]["x"], b["linearVelocity"]["y"], b["angularVelocity"]], axis=1).astype(np.float32)
    out["steps"] = np.int32(steps)
    return out


def main():
    from oracle_backend import OracleBackend
    here = os.path.dirname(os.path.abspath(__file__))
    for name, (build, steps, _) in CASES.items():
        w = api.World(OracleBackend(), scenes.world_def())
        data = record(w, build, steps)
        w.Destroy()
        np.savez_compressed(os.path.join(here, name + ".npz"), **data)
        print(name, "steps", steps, "contacts", len(data[f"s{steps - 1}_contact"]))


if __name__ == "__main__":
    main()
