"""ORACLE — TEST INFRASTRUCTURE ONLY.  **parity unpinned**.

Control flow of the reference's encode/decode path restated over ``oracle.prims`` /
``oracle.entropy``.  Each method cites the reference lines it follows:
  ColorModel.compress / .decompress           /root/reference/model/model.py:95-208
  AnalysisTransform / SparseSynthesisTransform /root/reference/model/transforms.py
  ScaledBlock / GenerativeUpBlock / ConditionEncoder /root/reference/model/blocks.py
  MeanScaleHyperprior_Map                      /root/reference/model/entropy_models.py:253-414
  MeanScaleHyperprior (two-model variant)      /root/reference/model/entropy_models.py:104-250,
                                               selected by model/model.py:22-27,131-135,196-200

``params`` maps the reference's state_dict names (e.g. ``g_a.pre_conv.0.kernel``) to numpy
arrays: ME kernel layout [K, Cin, Cout] ([Cin, Cout] for kernel_size 1), bias [1, Cout].
"""
import numpy as np

from . import prims
from .prims import OT, conv, conv_transpose, features_at, relu, leaky_relu
from .entropy import EntropyBottleneckOracle, GaussianConditionalOracle


class OracleCodec:
    def __init__(self, params, config):
        self.p = {k: np.asarray(v) for k, v in params.items()}
        self.cfg = config
        self.ablation_ga = config["g_a"].get("condition_ablation")
        self.src_cond_ga = bool(config["g_a"]["source_condition"])
        self.src_cond_gs = bool(config["g_s"]["source_condition"])
        self.dense = config["g_s"].get("dense", True)
        # ColorModel.update -> CompressionModel.update(force=True) (model/model.py:30-36)
        # "entropy_model_map" in the config (model/model.py:22-27): y and the latent quality map are coded by
        # two MeanScaleHyperprior models (no h_q) instead of one MeanScaleHyperprior_Map
        self.two_models = "entropy_model_map" in config
        self.ebs = {}
        for prefix in ("entropy_model", "entropy_model_map") if self.two_models else ("entropy_model",):
            eb = {k.split(".")[-1]: v for k, v in self.p.items() if k.startswith(prefix + ".entropy_bottleneck.")}
            self.ebs[prefix] = EntropyBottleneckOracle(eb)
        self.eb = self.ebs["entropy_model"]
        self.gc = GaussianConditionalOracle()
        self.trace = {}

    # ---------------------------------------------------------------- layer helpers
    def _w(self, name):
        return self.p[name + ".kernel"], self.p.get(name + ".bias")

    def _conv(self, x, name, k=3, stride=1):
        W, b = self._w(name)
        return conv(x, W, b, k, stride)

    def _convT(self, x, name, k, out_coords=None):
        W, b = self._w(name)
        return conv_transpose(x, W, b, k, 2, out_coords)

    def _seq(self, x, prefix, spec):
        """spec: list of ('c', idx, ksize) | ('r',) | ('l',) mirroring an nn.Sequential."""
        for item in spec:
            if item[0] == "c":
                x = self._conv(x, f"{prefix}.{item[1]}", item[2])
            elif item[0] == "r":
                x = x.with_feats(relu(x.F))
            elif item[0] == "l":
                x = x.with_feats(leaky_relu(x.F))
        return x

    # ---------------------------------------------------------------- blocks.py
    def scaled_block(self, x, condition, prefix):
        """ScaledBlock.forward (model/blocks.py:29-53); self.gdn is never called."""
        x_res = x.with_feats(x.F.copy())
        x = self._seq(x, prefix + ".conv_1", [("c", 0, 3), ("r",), ("c", 2, 3)])
        bg = features_at(condition, x.C)
        half = bg.shape[1] // 2
        beta, gamma = bg[:, :half], bg[:, half:]
        feats = x.F * beta + gamma  # two rounded fp32 ops, as eager torch (model/blocks.py:40)
        x = x.with_feats(feats.astype(np.float32))
        x = self._seq(x, prefix + ".conv_2", [("c", 0, 3), ("r",), ("c", 2, 3), ("r",)])
        return x.with_feats(x.F + features_at(x_res, x.C))

    def condition_encoder(self, Q, prefix):
        """ConditionEncoder.forward (model/blocks.py:235-251); conv_layers are never called."""
        Q = self._seq(Q, prefix + ".pre_conv", [("c", 0, 3), ("r",)])
        beta_gammas = []
        for i in range(3):
            Q = self._conv(Q, f"{prefix}.down_layers.{i}", 3, 2)
            if self.ablation_ga is None:
                bg = self._seq(Q, f"{prefix}.predict_layers.{i}",
                               [("c", 0, 3), ("r",), ("c", 2, 1), ("r",), ("c", 4, 3)])
            else:
                bg = Q
            beta_gammas.append(bg)
        return Q, beta_gammas

    def up_block(self, x, prefix, k):
        """GenerativeUpBlock.forward, predict=True (model/blocks.py:152-177)."""
        x = self._convT(x, prefix + ".conv", 3)
        conv2 = [("c", 0, 3), ("r",), ("c", 2, 3)]
        if self.dense:
            x = self._seq(x, prefix + ".conv_2", conv2)
            predictions = self._seq(x, prefix + ".occ_predict", conv2)
            mask = self.topk_prediction(predictions, k)
            up_coords = predictions.C[mask]
            x = self.prune_coords(x, up_coords)
        else:
            predictions = self._seq(x, prefix + ".occ_predict", conv2)
            mask = self.topk_prediction(predictions, k)
            up_coords = predictions.C[mask]
            x = self.prune_coords(x, up_coords)
            x = self._seq(x, prefix + ".conv_2", conv2)
        return x, predictions, up_coords

    def q_up_block(self, Q, prefix, coords):
        """GenerativeUpBlock.forward, predict=False (model/blocks.py:153,179-181)."""
        Q = self._convT(Q, prefix + ".conv", 3)
        return self.prune_coords(Q, coords)

    @staticmethod
    def topk_prediction(prediction, k):
        """_topk_prediction (model/blocks.py:130-150): per batch id, top-k of channel 0."""
        mask = np.zeros(prediction.n, dtype=bool)
        for j, b in enumerate(np.unique(prediction.C[:, 0])):
            sel = np.nonzero(prediction.C[:, 0] == b)[0]
            mask[sel[prims.topk_mask(prediction.F[sel, 0], int(k[int(b)]))]] = True
        return mask

    @staticmethod
    def prune_coords(x, occupied):
        """_prune_coords (model/blocks.py:101-128): torch.isin on int64 keys + MinkowskiPruning
        (keeps rows in order)."""
        mask = np.isin(x.keys, prims.pack_keys(occupied))
        return OT(x.C[mask], x.F[mask], x.ts, _sorted=True)

    # ---------------------------------------------------------------- transforms.py
    @staticmethod
    def count_per_batch(x):
        """AnalysisTransform.count_per_batch (model/transforms.py:65-71)."""
        return [int((x.C[:, 0] == b).sum()) for b in np.unique(x.C[:, 0])]

    def g_a(self, x, Q):
        """AnalysisTransform.forward (model/transforms.py:75-128)."""
        k = [self.count_per_batch(x)]
        if self.src_cond_ga:
            Q_plus = self._seq(x, "g_a.cond_conv", [("c", 0, 3), ("r",), ("c", 2, 3)])
            Q = OT(Q.C, Q.F + features_at(Q_plus, Q.C), 1, _sorted=True)
        Q, bgs = self.condition_encoder(Q, "g_a.condition_encoder")
        x = self._seq(x, "g_a.pre_conv", [("c", 0, 3), ("r",)])
        x = self._conv(x, "g_a.down_1", 3, 2)
        x = self.scaled_block(x, bgs[0], "g_a.scale_1")
        k.append(self.count_per_batch(x))
        x = self._conv(x, "g_a.down_2", 3, 2)
        x = self.scaled_block(x, bgs[1], "g_a.scale_2")
        k.append(self.count_per_batch(x))
        x = self._conv(x, "g_a.down_3", 3, 2)
        x = self.scaled_block(x, bgs[2], "g_a.scale_3")
        x = self._conv(x, "g_a.post_conv", 3)
        Q = OT(x.C, features_at(Q, x.C), x.ts, _sorted=True)
        k.reverse()
        return x, Q, k

    def g_s(self, x, Q, k):
        """SparseSynthesisTransform.forward, coords=None (model/transforms.py:242-304)."""
        if self.src_cond_gs:
            Q_plus = self._seq(x, "g_s.cond_conv", [("c", 0, 3), ("r",), ("c", 2, 3)])
            Q = OT(Q.C, Q.F + features_at(Q_plus, Q.C), Q.ts, _sorted=True)
        x = self._seq(x, "g_s.pre_conv", [("c", 0, 3), ("r",)])
        Q = self._seq(Q, "g_s.q_pre_conv", [("c", 0, 3), ("r",), ("c", 2, 1), ("r",), ("c", 4, 3)])
        pred = [("c", 0, 3), ("r",), ("c", 2, 3), ("r",), ("c", 4, 3)]
        for s in (1, 2, 3):
            beta_gamma = self._seq(Q, f"g_s.q_predict_{s}", pred)
            x = self.scaled_block(x, beta_gamma, f"g_s.scale_{s}")
            x, predictions, up_coords = self.up_block(x, f"g_s.up_{s}", k[s - 1])
            Q = self.q_up_block(Q, f"g_s.q_up_{s}", up_coords)
            self.trace[f"up_coords_{s}"] = up_coords
            self.trace[f"logits_{s}"] = predictions.F[:, 0].copy()
            self.trace[f"cand_{s}"] = predictions.C
        x = self._seq(x, "g_s.post_conv", [("c", 0, 3), ("r",), ("c", 2, 3), ("r",), ("c", 4, 3)])
        return x

    # ---------------------------------------------------------------- entropy_models.py
    def h_a(self, y, model="entropy_model"):
        P = model + ".h_a"
        z = self._conv(y, P + ".0", 3)
        z = z.with_feats(leaky_relu(z.F))
        z = self._conv(z, P + ".2", 3, 2)
        z = self._conv(z, P + ".3", 3)
        z = z.with_feats(leaky_relu(z.F))
        z = self._conv(z, P + ".5", 3, 2)
        return self._conv(z, P + ".6", 3)

    def h_s(self, z_hat, model="entropy_model"):
        """h_s (model/entropy_models.py:284-294, :135-145); the Sorted* shims (:12-102) only permute rows
        into canonical order, which every oracle tensor already has."""
        P = model + ".h_s"
        g = self._conv(z_hat, P + ".0", 3)
        g = self._convT(g, P + ".1", 2)
        g = g.with_feats(leaky_relu(g.F))
        g = self._conv(g, P + ".3", 3)
        g = self._convT(g, P + ".4", 2)
        g = g.with_feats(leaky_relu(g.F))
        return self._conv(g, P + ".6", 3)

    def h_q(self, z_hat):
        """h_q (model/entropy_models.py:296-306): z_hat has a fresh coordinate manager, so both
        MinkowskiConvolutionTranspose layers generate the 27-dilated map."""
        P = "entropy_model.h_q"
        q = self._conv(z_hat, P + ".0", 3)
        q = self._convT(q, P + ".1", 3)
        q = q.with_feats(relu(q.F))
        q = self._conv(q, P + ".3", 3)
        q = self._convT(q, P + ".4", 3)
        q = q.with_feats(relu(q.F))
        return self._conv(q, P + ".6", 3)

    def _gaussian_params(self, z_hat, y_coords, model="entropy_model"):
        gp = features_at(self.h_s(z_hat, model), y_coords)
        half = gp.shape[1] // 2
        return gp[:, :half].T.copy(), gp[:, half:].T.copy()  # scales, means as [C, N]

    def entropy_compress(self, y, model="entropy_model"):
        """MeanScaleHyperprior_Map.compress (model/entropy_models.py:341-381) and MeanScaleHyperprior.compress
        (:179-218): the same statements, the second class only lacks h_q."""
        eb = self.ebs[model]
        z = self.h_a(y, model)  # y, z already canonical (sort_tensor, :346-347 / :184-185)
        shape = [z.n]
        z_string = eb.compress(z.F.T)
        z_hat_feats = eb.decompress(z_string, shape[0])  # (:353 / :191) decode again to get z_hat
        z_hat = OT(z.C, z_hat_feats.T, 32, _sorted=True)
        scales, means = self._gaussian_params(z_hat, y.C, model)
        indexes = self.gc.build_indexes(scales)
        y_string = self.gc.compress(y.F.T, indexes, means)
        sfx = "" if model == "entropy_model" else "_map"
        self.trace.update({"z" + sfx: z, "z_hat" + sfx: z_hat, "scales" + sfx: scales, "means" + sfx: means,
                           "indexes" + sfx: indexes, "y_symbols" + sfx: self.gc.quantize(y.F.T, means),
                           "z_symbols" + sfx: np.rint(z.F.T - eb.medians()[:, None]).astype(np.int32)})
        return [y.C, z.C], [[y_string], [z_string]], shape

    def entropy_decompress(self, points, strings, shape, model="entropy_model", with_h_q=True):
        """MeanScaleHyperprior_Map.decompress (model/entropy_models.py:384-414) -> (y_hat, Q_hat);
        MeanScaleHyperprior.decompress (:221-250, with_h_q=False) -> (y_hat, None)."""
        y_points = prims.unpack_keys(np.sort(prims.pack_keys(points[0])))  # sort_points (:389-390 / :226-227)
        z_points = prims.unpack_keys(np.sort(prims.pack_keys(points[1])))
        z_hat = OT(z_points, self.ebs[model].decompress(strings[1][0], shape[0]).T, 32, _sorted=True)
        Q_hat = self.h_q(z_hat) if with_h_q else None
        scales, means = self._gaussian_params(z_hat, y_points, model)
        indexes = self.gc.build_indexes(scales)
        y_hat = OT(y_points, self.gc.decompress(strings[0][0], indexes, means).T, 8, _sorted=True)
        return y_hat, Q_hat

    def entropy_forward(self, y):
        """MeanScaleHyperprior_Map.forward, eval mode (model/entropy_models.py:309-337)."""
        z = self.h_a(y)
        z_hat_f, z_lik = self.eb.forward(z.F.T)
        z_hat = OT(z.C, z_hat_f.T, 32, _sorted=True)
        Q_hat = self.h_q(z_hat)
        scales, means = self._gaussian_params(z_hat, y.C)
        y_hat_f, y_lik = self.gc.forward(y.F.T, scales, means)
        return OT(y.C, y_hat_f.T, 8, _sorted=True), Q_hat, (y_lik, z_lik)

    # ---------------------------------------------------------------- model.py
    def compress(self, x, Q_coords, Q_feats):
        """ColorModel.compress, path=None (model/model.py:95-147).
        x: [N,6] float (xyz, rgb); Q: coordinates [N,4] + features [N,2] (utils.py:436-445)."""
        N = x.shape[0]
        points = np.concatenate([np.zeros((N, 1), np.float32), x[:, :3]], axis=1).astype(np.int32)
        feats = np.concatenate([np.ones((N, 1), np.float32), x[:, 3:6]], axis=1)
        inp = OT(points, feats, 1)
        Q = OT(Q_coords, Q_feats, 1)
        y, Q, k = self.g_a(inp, Q)
        self.trace.update(y=y, Q_latent=Q)
        if not self.two_models:
            points, strings, shape = self.entropy_compress(y)
        else:  # model/model.py:131-135
            points, y_strings, y_shape = self.entropy_compress(y, "entropy_model")
            _, Q_strings, Q_shape = self.entropy_compress(Q, "entropy_model_map")
            strings = [y_strings, Q_strings]
            shape = [y_shape, Q_shape]
        return strings, shape, k, y.C

    def decompress(self, coordinates, strings, shape, k):
        """ColorModel.decompress, path=None (model/model.py:152-208)."""
        lat = OT(coordinates, np.ones((coordinates.shape[0], 1), np.float32), 8)
        W = self.p["g_s.down_conv.kernel"]
        lat = conv(conv(lat, W, None, 3, 2), W, None, 3, 2)  # only .C is used (:188-191)
        if not self.two_models:
            y_hat, Q_hat = self.entropy_decompress([coordinates, lat.C], strings, shape)
        else:  # model/model.py:196-200
            points = [coordinates, lat.C]
            y_hat, _ = self.entropy_decompress(points, strings[0], shape[0], "entropy_model", with_h_q=False)
            Q_hat, _ = self.entropy_decompress(points, strings[1], shape[1], "entropy_model_map", with_h_q=False)
        self.trace.update(y_hat=y_hat, Q_hat=Q_hat)
        x_hat = self.g_s(y_hat, Q_hat, k)
        feats = np.clip(np.rint(x_hat.F * np.float32(255)), 0.0, 255.0).astype(np.float32) / np.float32(255)
        return np.concatenate([x_hat.C[:, 1:4].astype(np.float32), feats], axis=1)
