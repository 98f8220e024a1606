#!/usr/bin/env python
"""Generates tests/golden/*.npz by running the REFERENCE'S OWN model code.

  python tests/golden/make_golden.py          (in the builder container: needs /root/reference)

/root/reference/model/{model,transforms,blocks,entropy_models}.py and utils.py are imported
unmodified; the third-party packages they need and that cannot be installed offline
(MinkowskiEngine, compressai, open3d, bitstream) are replaced by the CPU stand-ins of
oracle/ref_shims.py, which implement SURVEY.md Appendix A/B over the oracle primitives.  The
fixtures are therefore outputs of ColorModel.compress / ColorModel.decompress AS WRITTEN IN THE
REFERENCE (model/model.py:95-208), with the reference's configs/*.yaml, on deterministic synthetic
weights and inputs.  They pin:  the oracle's control-flow restatement (oracle/codec.py), the
product's fused plan (codec.py) and the state_dict naming, against the reference's code.
They do not pin MinkowskiEngine / CompressAI internals (absent; "parity unpinned" there).
"""
import os
import sys

import numpy as np
import torch
import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"

CASES = [  # name, yaml, product config name, shell radius, seed, (q_g, q_a), non-uniform map or None
    ("ours_r14", "configs/Ours.yaml", "OURS_CONFIG", 14.0, 0, (0.5, 0.5), None),
    ("ours_r11_q", "configs/Ours.yaml", "OURS_CONFIG", 11.0, 3, (0.2, 0.9), None),
    ("ablation_r12", "configs/Ablation_NoCondition_Convolution.yaml", "ABLATION_CONFIG", 12.0, 1, (0.75, 0.25), None),
    # view-dependent gradient along y and region of interest beyond x = 64 (evaluate_view_dep.py:209-216,252-258)
    ("ours_r13_view", "configs/Ours.yaml", "OURS_CONFIG", 13.0, 4, (0.8, 1.0), ("view", 1, 52.0, 76.0)),
    ("ours_r12_roi", "configs/Ours.yaml", "OURS_CONFIG", 12.0, 5, (1.0, 0.6), ("roi", 0, 64.0)),
    # ColorModel.__init__ with an "entropy_model_map" section (model/model.py:22-27): two MeanScaleHyperprior models
    # (model/entropy_models.py:104-250) code y and the latent quality map separately.  No shipped YAML has the
    # section; it is added to Ours.yaml's `model:` here and stored in the fixture as JSON.
    ("twohyper_r13", "configs/Ours.yaml", "OURS_CONFIG", 13.0, 6, (0.4, 0.8), None),
]
EXTRA = {"twohyper_r13": {"entropy_model": {"C_bottleneck": 128, "C_hyper_bottleneck": 128},
                          "entropy_model_map": {"C_bottleneck": 2, "C_hyper_bottleneck": 8}}}


def quality_features(x, q, kind):
    """Per-point [q_g, q_a] the way the reference's scripts build them (numpy, float64 scores)."""
    pts = x.numpy()[:, :3].astype(np.float64)
    if kind is None:
        return None
    if kind[0] == "view":
        _, axis, lo, hi = kind
        scores = np.clip((pts[:, axis] - lo) / (hi - lo), 0, 1)
    else:
        _, axis, plane = kind
        scores = np.where(pts[:, axis] < plane, 0, 1)
    q_map_g = (q[0] * scores).reshape(-1, 1)
    q_map_a = (q[1] * scores).reshape(-1, 1)
    return torch.cat([torch.tensor(q_map_g), torch.tensor(q_map_a)], dim=1).type(torch.float32)  # utils.py:442


def string_fields(name, cfg, strings, shape):
    """One hyperprior: strings = [[y], [z]], shape = [N_z].  Two (model/model.py:131-135): strings =
    [[[y], [z]], [[y_map], [z_map]]], shape = [[N_z], [N_z_map]]; the config goes along as JSON."""
    import json
    u8 = lambda b: np.frombuffer(b, np.uint8)
    if "entropy_model_map" not in cfg:
        return dict(y_string=u8(strings[0][0]), z_string=u8(strings[1][0]), shape=np.array(shape, np.int64))
    return dict(y_string=u8(strings[0][0][0]), z_string=u8(strings[0][1][0]), y_map_string=u8(strings[1][0][0]),
                z_map_string=u8(strings[1][1][0]), shape=np.array([shape[0][0], shape[1][0]], np.int64),
                config_json=np.array(json.dumps(cfg)))


def main():
    import pcc_b200
    from oracle import ref_shims
    for name, yml, cfg_name, radius, seed, q, kind in CASES:
        cfg = yaml.safe_load(open(os.path.join(REF, yml)))["model"]
        assert cfg == getattr(pcc_b200, cfg_name), f"{yml} differs from pcc_b200.{cfg_name}"
        cfg.update(EXTRA.get(name, {}))
        ref_model = ref_shims.load_reference_model(cfg, REF)
        # deterministic synthetic weights, taken through the state_dict so key names are pinned
        ours = pcc_b200.ColorModel(cfg)
        pcc_b200.init_synthetic(ours, seed)
        sd = {k: v for k, v in ours.state_dict().items()}
        missing, unexpected = ref_model.load_state_dict(sd, strict=False)
        assert not missing, missing          # every parameter of the reference's module tree is supplied
        assert all(("_offset" in u or "_quantized_cdf" in u or "_cdf_length" in u) for u in unexpected), unexpected
        ref_model.eval()
        ref_model.update()
        x = pcc_b200.synthetic.shell_frame(radius, center=[64, 64, 64], grid=128, seed=seed)
        qc, qf = pcc_b200.synthetic.uniform_q_map(x, *q)
        if kind is not None:
            qf = quality_features(x, q, kind)
        import MinkowskiEngine as ME  # the stand-in
        with torch.no_grad():
            Q = ME.SparseTensor(coordinates=qc, features=qf)     # utils.py:436-445
            strings, shape, k, coordinates = ref_model.compress(x, Q)
            out = ref_model.decompress(coordinates=coordinates.clone(), strings=strings, shape=shape, k=k)
        # canonical order for comparison (ME's row order is unspecified)
        out = out.numpy()
        key = (out[:, 0].astype(np.int64) << 40) | (out[:, 1].astype(np.int64) << 20) | out[:, 2].astype(np.int64)
        out = out[np.argsort(key)]
        coords = coordinates.numpy().astype(np.int32)
        ck = (coords[:, 1].astype(np.int64) << 40) | (coords[:, 2].astype(np.int64) << 20) | coords[:, 3].astype(np.int64)
        coords = coords[np.argsort(ck)]
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"),
            config=cfg_name, radius=radius, seed=seed, q=np.array(q, np.float32),
            x=x.numpy().astype(np.float32), qf=qf.numpy().astype(np.float32),
            qkind=np.array([] if kind is None else [{"view": 0.0, "roi": 1.0}[kind[0]]] + [float(v) for v in kind[1:]]),
            k=np.array([int(kk[0]) for kk in k], np.int64), coordinates=coords, out=out.astype(np.float32),
            **string_fields(name, cfg, strings, shape))
        print(f"{name}: {x.shape[0]} points, string bytes {[[len(b[0]) for b in s] if isinstance(s[0], list) else len(s[0]) for s in strings]}, "
              f"shape {shape}, k {k}, latent voxels {coords.shape[0]}, decoded {out.shape}")
        for m in [m for m in sys.modules if m == "model" or m.startswith("model.") or m == "utils"]:
            del sys.modules[m]


if __name__ == "__main__":
    main()
