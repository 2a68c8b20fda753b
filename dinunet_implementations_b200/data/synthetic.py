"""Synthetic simulator datasets in the reference's on-disk layout.

The reference ships one toy dataset (``datasets/test_fsl``: 5 sites with 73 / 50 / 100 / 80 / 120
subjects, one ``subjectN_aseg_stats.txt`` per subject — header line + 66 ``name<TAB>value`` rows —
and a ``site<k>_Covariate.csv`` with columns ``freesurferfile,isControl,age``) and only the
``inputspec.json`` of the ICA task (``datasets/icalstm/inputspec.json``: 2 sites, data file
``HCP_AllData_sess1.npz``, labels ``S2_<k>.csv``; the arrays themselves are not shipped).  Real
data cannot be fetched here, so this module *generates* datasets with the same directory layout,
file formats, field names and site sizes, with a planted class signal so that training has
something to learn:

    <root>/inputspec.json
    <root>/input/local<i>/simulatorRun/{subject*_aseg_stats.txt, site<i+1>_Covariate.csv}     (FS)
    <root>/input/local<i>/simulatorRun/{<data_file>.npz, S2_<i+1>.csv}                         (ICA)

``python -m dinunet_implementations_b200.data.synthetic --task fs  --out datasets/test_fsl``
``python -m dinunet_implementations_b200.data.synthetic --task ica --out datasets/icalstm``
"""
from __future__ import annotations

import argparse
import csv
import json
import os

import numpy as np

FS_SITE_SIZES = (73, 50, 100, 80, 120)       # reference datasets/test_fsl site sizes
FS_NUM_REGIONS = 66

# Segmentation labels in the order an aseg stats table lists them (FreeSurfer's public atlas names).
_ASEG_HEAD = ["Lateral-Ventricle", "Inf-Lat-Vent", "Cerebellum-White-Matter", "Cerebellum-Cortex",
              "Thalamus-Proper", "Caudate", "Putamen", "Pallidum"]
_ASEG_MID = ["3rd-Ventricle", "4th-Ventricle", "Brain-Stem"]
_ASEG_TAIL = ["Hippocampus", "Amygdala"]
_ASEG_REST = ["CSF", "Optic-Chiasm", "CC_Posterior", "CC_Mid_Posterior", "CC_Central", "CC_Mid_Anterior",
              "CC_Anterior", "BrainSegVol", "BrainSegVolNotVent", "BrainSegVolNotVentSurf", "lhCortexVol",
              "rhCortexVol", "CortexVol", "lhCerebralWhiteMatterVol", "rhCerebralWhiteMatterVol",
              "CerebralWhiteMatterVol", "SubCortGrayVol", "TotalGrayVol", "SupraTentorialVol",
              "SupraTentorialVolNotVent", "SupraTentorialVolNotVentVox", "MaskVol", "BrainSegVol-to-eTIV",
              "MaskVol-to-eTIV", "lhSurfaceHoles", "rhSurfaceHoles", "SurfaceHoles", "EstimatedTotalIntraCranialVol"]


def aseg_region_names(n=FS_NUM_REGIONS):
    names = [f"Left-{r}" for r in _ASEG_HEAD] + _ASEG_MID + [f"Left-{r}" for r in _ASEG_TAIL]
    names += ["Left-Accumbens-area", "Left-VentralDC", "Left-vessel", "Left-choroid-plexus"]
    names += [f"Right-{r}" for r in _ASEG_HEAD + _ASEG_TAIL]
    names += ["Right-Accumbens-area", "Right-VentralDC", "Right-vessel", "Right-choroid-plexus"]
    names += ["5th-Ventricle", "WM-hypointensities", "non-WM-hypointensities"] + _ASEG_REST
    while len(names) < n:
        names.append(f"Region-{len(names)}")
    return names[:n]


def _wrap(d):
    return {k: {"value": v} for k, v in d.items()}


def make_fs_dataset(root, site_sizes=FS_SITE_SIZES, num_regions=FS_NUM_REGIONS, seed=0, effect=0.35,
                    extra_inputs=None):
    """Write a FreeSurfer-volumes dataset; returns the list of per-site input dicts.

    Each region has a log-normal population volume; controls and patients differ by a fixed random
    direction of relative size ``effect`` on a third of the regions, plus an age trend and noise.
    """
    rng = np.random.default_rng(seed)
    names = aseg_region_names(num_regions)
    base = np.exp(rng.uniform(np.log(300.0), np.log(60000.0), size=num_regions))
    direction = rng.choice([-1.0, 0.0, 1.0], size=num_regions, p=[1 / 6, 2 / 3, 1 / 6])
    age_slope = rng.normal(0.0, 0.002, size=num_regions)
    spec = []
    for s, n in enumerate(site_sizes):
        d = os.path.join(root, "input", f"local{s}", "simulatorRun")
        os.makedirs(d, exist_ok=True)
        site_gain = 1.0 + rng.normal(0.0, 0.03)            # scanner / site effect
        rows = []
        for i in range(int(n)):
            control = bool(rng.integers(0, 2))
            age = float(np.round(rng.uniform(18, 80)))
            rel = 1.0 + effect * direction * (0.5 if control else -0.5) + age_slope * (age - 45.0)
            vol = base * rel * site_gain * np.exp(rng.normal(0.0, 0.12, size=num_regions))
            fname = f"subject{i}_aseg_stats.txt"
            with open(os.path.join(d, fname), "w") as f:
                f.write(f"Measure:volume\tsubject{i}\n")
                for nm, v in zip(names, vol):
                    f.write(f"{nm}\t{v:.2f}\n")
            rows.append((fname, str(control), age))
        cov = f"site{s + 1}_Covariate.csv"
        with open(os.path.join(d, cov), "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["freesurferfile", "isControl", "age"])
            w.writerows(rows)
        inp = {"labels_file": cov, "data_column": "freesurferfile", "labels_column": "isControl", "mode": "train",
               "gpus": [], "input_size": int(num_regions), "hidden_sizes": [256, 128, 64, 32], "num_class": 2,
               "num_workers": 2, "learning_rate": 0.001}
        inp.update(extra_inputs or {})
        spec.append(inp)
    with open(os.path.join(root, "inputspec.json"), "w") as f:
        json.dump([_wrap(i) for i in spec], f, indent=2)
    return spec


def make_ica_dataset(root, num_sites=2, subjects_per_site=48, num_components=100, temporal_size=980,
                     window_size=10, window_stride=10, input_size=256, hidden_size=348, seed=0,
                     data_file="HCP_AllData_sess1.npz", effect=0.6, gpus_per_site=True, extra_inputs=None):
    """Write an ICA time-course dataset (``[subjects, components, time]`` float32 in an ``.npz``
    plus a two-column ``index,label`` CSV per site); returns the per-site input dicts.

    Time courses are AR(1) noise mixed through a random component-coupling matrix; class 1 adds a
    slow oscillation on a fixed subset of components, so the signal is temporal, not a mean shift.
    """
    rng = np.random.default_rng(seed)
    mix = np.eye(num_components) + 0.15 * rng.normal(size=(num_components, num_components)) / np.sqrt(num_components)
    marked = rng.choice(num_components, size=max(1, num_components // 8), replace=False)
    phase = rng.uniform(0, 2 * np.pi, size=marked.size)
    t = np.arange(temporal_size)
    spec = []
    for s in range(num_sites):
        d = os.path.join(root, "input", f"local{s}", "simulatorRun")
        os.makedirs(d, exist_ok=True)
        n = int(subjects_per_site if np.isscalar(subjects_per_site) else subjects_per_site[s])
        labels = rng.integers(0, 2, size=n)
        e = rng.normal(size=(n, num_components, temporal_size)).astype(np.float32)
        x = np.empty_like(e)
        x[:, :, 0] = e[:, :, 0]
        for k in range(1, temporal_size):                   # AR(1), rho = 0.7
            x[:, :, k] = 0.7 * x[:, :, k - 1] + e[:, :, k]
        x = np.einsum("ij,njt->nit", mix.astype(np.float32), x)
        osc = np.sin(2 * np.pi * t[None, :] / 49.0 + phase[:, None]).astype(np.float32)
        x[np.ix_(np.nonzero(labels)[0], marked)] += effect * osc[None]
        np.savez(os.path.join(d, data_file), data=x.astype(np.float32))
        lab = f"S2_{s + 1}.csv"
        with open(os.path.join(d, lab), "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["index", "label"])
            w.writerows([[i, int(y)] for i, y in enumerate(labels)])
        inp = {"mode": "train", "gpus": [s] if gpus_per_site else [], "num_class": 2, "num_workers": 2,
               "learning_rate": 0.001, "data_file": data_file, "labels_file": lab, "input_size": int(input_size),
               "hidden_size": int(hidden_size), "window_size": int(window_size), "window_stride": int(window_stride),
               "temporal_size": int(temporal_size), "num_components": int(num_components)}
        inp.update(extra_inputs or {})
        spec.append(inp)
    with open(os.path.join(root, "inputspec.json"), "w") as f:
        json.dump([_wrap(i) for i in spec], f, indent=2)
    return spec


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--task", choices=["fs", "ica"], required=True)
    ap.add_argument("--out", required=True)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--sites", type=int, default=None)
    ap.add_argument("--subjects", type=int, default=None, help="subjects per site (default: reference site sizes)")
    a = ap.parse_args(argv)
    if a.task == "fs":
        sizes = FS_SITE_SIZES if a.sites is None else tuple((a.subjects or FS_SITE_SIZES[i % 5]) for i in range(a.sites))
        if a.subjects and a.sites is None:
            sizes = tuple(a.subjects for _ in FS_SITE_SIZES)
        make_fs_dataset(a.out, sizes, seed=a.seed)
    else:
        make_ica_dataset(a.out, num_sites=a.sites or 2, subjects_per_site=a.subjects or 48, seed=a.seed)
    print(f"wrote {a.task} dataset to {a.out}")


if __name__ == "__main__":
    main()
