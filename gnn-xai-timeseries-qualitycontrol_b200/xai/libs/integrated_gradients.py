"""Drop-in for the attribution core of the reference's xai/libs/integrated_gradients.py
(IntegratedGradientsExplainer): zero baseline, m_steps+1 interpolation points, trapezoid, scaling by
the input, thresholded class, per-sample unwrap and the six .npy files per sample that
xai/libs/integrated_gradients_analyser.py consumes.  The 51/101 forward+backward passes per batch
run inside libgnnqc.so (gnnqc_ig_cml); nothing here touches TensorFlow.

Out of scope, as in SURVEY.md section 2: TFRecord loading, logging, every plot.

Reference lines restated by each method are cited in the docstrings.
"""
import os
from pathlib import Path

import numpy as np
import torch

from gnnqc import dp


def split_work(items, workerid, n_worker):
    """Reference :432-448: item i belongs to worker i % n_worker (SLURM array sharding; here one
    worker per GPU / rank)."""
    return [x for i, x in enumerate(items) if i % n_worker == workerid]


class IntegratedGradientsExplainer:
    def __init__(self, model, xai_config, preproc_config):
        """model: libs.create_model.GCNClassifier (CML).  xai_config / preproc_config: attribute
        dicts with the keys of xai_config_20240318.yml / preprocessing_config_20240318.yml."""
        if model.ds_type != "cml" or not model.native.has_gcn:
            raise NotImplementedError("integrated gradients are built for the CML GCN model")
        self.model = model
        self.xai_config = xai_config
        self.preproc_config = preproc_config
        self.ig = xai_config.integrated_gradients
        if self.ig.baseline != "zero":
            raise NotImplementedError("only the zero baseline is implemented (as in the reference, :898-917)")
        self.rank, self.world_size = dp.world()

    # ---- core (:1133-1234) -------------------------------------------------------------------
    def get_gradients_single_batch(self, batch):
        """batch: offsets-layout dict (features, anom, edges, node_off, edge_off).  Returns a dict
        with gradients_features [Ntot,2], gradients_anom_ts [B,T,2], predictions [B,1],
        predicted class [B,1] (all numpy)."""
        dev = torch.device("cuda", self.model.device)
        b = {k: (v.to(dev) if isinstance(v, torch.Tensor) else v) for k, v in batch.items()}
        af, aa, p = self.model.native.ig_cml(b["features"], b["anom"], b["edges"], b["node_off"].to(torch.int32), b["edge_off"],
                                             m_steps=int(self.ig.m_steps), n_nodes_total=batch.get("n_nodes_total"),
                                             scale_by_input=bool(self.ig.scale_gradients))    # :1185-1191 only when configured
        mode = self.ig.negative_values             # :1193-1207
        if mode == "abs":
            af, aa = af.abs(), aa.abs()
        elif mode == "clip":
            af, aa = af.clamp(min=0), aa.clamp(min=0)
        elif mode != "keep":
            raise ValueError("integrated_gradients.negative_values must be 'abs', 'clip' or 'keep', got %r" % (mode,))
        pred = p.reshape(-1, 1).cpu().numpy()
        return {
            "gradients_features": af.cpu().numpy(), "gradients_anom_ts": aa.cpu().numpy(), "predictions": pred,
            "predicted_class": np.where(pred > float(self.ig.threshold), 1, 0),     # :1229-1234
        }

    # ---- unwrap (:1017-1042) ------------------------------------------------------------------
    @staticmethod
    def unwrap_features(index, rows, node_off, T):
        """Rows of sample `index` ((t, n) order) -> [N_b, T, F]: the reference's a[i::n_neighbors]."""
        n0, n1 = int(node_off[index]), int(node_off[index + 1])
        a = np.asarray(rows[n0 * T: n1 * T])
        return a.reshape(T, n1 - n0, a.shape[1]).transpose(1, 0, 2)

    # ---- file naming (:273-384) ------------------------------------------------------------------
    def _sample_stem(self, sensor_id, anomaly_date, true, pred):
        return "_".join([self.xai_config.project, self.preproc_config.ds_type, self.ig.dataset, str(sensor_id), str(anomaly_date),
                         str(int(true)), str(int(pred))])

    def _sample_dir(self, sensor_id, anomaly_date, true, pred):
        return Path(self.xai_config.output_dir, "integrated_gradients", self.xai_config.project, self.preproc_config.ds_type,
                    self.ig.dataset, str(sensor_id), "_".join([str(sensor_id), str(anomaly_date), str(int(true)), str(int(pred))]))

    # ---- per-sample files (:1248-1370) ---------------------------------------------------------------
    def unwrap_and_save(self, batch, result, sensor_ids, anomaly_dates, anomaly_flags_true):
        """Writes, for every sample whose confusion class is in which_samples, the six arrays the
        analyser reads: gradients_features_unwrapped, gradients_anom_ts_unwrapped, features_unwrapped,
        anom_ts_unwrapped, predictions_unwrapped, anomaly_flag_true_unwrapped.  Returns the paths."""
        T = batch["anom"].shape[1]
        node_off = np.asarray(batch["node_off"].cpu())
        feats = np.asarray(batch["features"].cpu())
        anom = np.asarray(batch["anom"].cpu())
        wanted = set(self.ig.which_samples)
        written = []
        for i in range(anom.shape[0]):
            true, pred = int(anomaly_flags_true[i]), int(result["predicted_class"][i, 0])
            kind = ("T" if true == pred else "F") + ("P" if pred == 1 else "N")
            if kind not in wanted:
                continue
            out_dir = self._sample_dir(sensor_ids[i], anomaly_dates[i], true, pred)
            os.makedirs(out_dir, exist_ok=True)
            stem = self._sample_stem(sensor_ids[i], anomaly_dates[i], true, pred)
            arrays = {
                "gradients_features_unwrapped_": self.unwrap_features(i, result["gradients_features"], node_off, T),
                "gradients_anom_ts_unwrapped_": result["gradients_anom_ts"][i],
                "features_unwrapped_": self.unwrap_features(i, feats, node_off, T),
                "anom_ts_unwrapped_": anom[i],
                "predictions_unwrapped_": result["predictions"][i],
                "anomaly_flag_true_unwrapped_": np.asarray(true),
            }
            for prefix, arr in arrays.items():
                path = out_dir / (prefix + stem + ".npy")
                np.save(path, arr)
                written.append(path)
        return written

    def get_gradients(self, batches, meta=None):
        """Runs every batch this rank owns (reference :1093-1131; batches are dealt round-robin over
        ranks like the reference deals files over SLURM array tasks)."""
        results = []
        for i, batch in enumerate(batches):
            if i % self.world_size != self.rank:
                continue
            res = self.get_gradients_single_batch(batch)
            if meta is not None:
                res["files"] = self.unwrap_and_save(batch, res, *meta[i])
            results.append(res)
        return results
