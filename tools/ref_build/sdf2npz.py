#!/usr/bin/env python3
"""sdf2npz.py RUNDIR OUTDIR NAME: rewrite the SDF dumps 0000.sdf, 0001.sdf of an EPOCH run as
NAME_0.npz, NAME_1.npz (plain numpy, readable without the sdf module).  Needs the `sdf` Python
module built from the SDF submodule (SDF/utilities/build)."""
import glob
import os
import sys

import numpy as np
import sdf


def convert(path):
    d = sdf.read(path, dict=True)
    out = {"time": np.float64(d["Header"]["time"]), "step": np.int64(d["Header"]["step"])}
    for k, name in (("Electric Field/E", "e"), ("Magnetic Field/B", "b"), ("Current/J", "j")):
        for c in "xyz":
            key = k + c
            if key in d:
                out[name + c] = np.asfortranarray(d[key].data).astype(np.float64)
    species = sorted(k.split("/")[-1] for k in d if k.startswith("Particles/Px/"))
    out["species"] = np.array(species)
    for sp in species:
        grid = d["Grid/Particles/" + sp].data          # tuple of position arrays
        for a, arr in zip("xyz", grid):
            out[f"{sp}_{a}"] = np.asarray(arr, dtype=np.float64)
        for q in ("Px", "Py", "Pz"):
            out[f"{sp}_{q.lower()}"] = np.asarray(d[f"Particles/{q}/{sp}"].data, dtype=np.float64)
        out[f"{sp}_w"] = np.asarray(d[f"Particles/Weight/{sp}"].data, dtype=np.float64)
    g = d["Grid/Grid"].data                            # cell edges
    out["edges"] = np.array([np.asarray(a)[[0, -1]] for a in g], dtype=np.float64)
    out["cells"] = np.array([len(a) - 1 for a in g], dtype=np.int64)
    return out


def main():
    run, outdir, name = sys.argv[1:4]
    files = sorted(glob.glob(os.path.join(run, "[0-9][0-9][0-9][0-9].sdf")))
    assert len(files) >= 2, files
    for i, f in enumerate(files[:2]):
        np.savez(os.path.join(outdir, f"{name}_{i}.npz"), **convert(f))
        print("wrote", os.path.join(outdir, f"{name}_{i}.npz"))


if __name__ == "__main__":
    main()
