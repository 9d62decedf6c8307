"""LDA-CGS application (``edu.iu.lda``, ``ml/java/src/main/java/edu/iu/lda/LDAMPCollectiveMapper.java``).

Documents ``docName wid wid ...`` are parsed per worker (``DocStore.java:157-168``), the vocabulary
size is agreed by ``allreduce`` of the max word id (``LDAUtil.java:209``), then the rotation engine
runs ``num_iterations`` of sampling.  ``<min/max training percentage>`` keep their meaning: the
per-pass time budget is halved / doubled so that the fraction of tokens sampled per iteration stays
inside ``[min, max]`` % (``:477-560``; 100/100 = full passes, no timer).  Every ``print_interval``
iterations (reference: 10) the log-likelihood is computed; the word-topic model is written as
``<work>/model/tmp_word_model/<iter>/<rank>`` when ``print_model`` and the last likelihood goes to
``<work>/model/evaluation`` (``:593-628``, ``:754-774``).
"""
from __future__ import annotations

import os
import time
from typing import List, Optional

import numpy as np
import torch

from ..io import formats as F
from ..mapper import CollectiveMapper, Context, KeyValReader, LOG
from ..models.lda import LDAEngine
from ..partition import DoubleArrMax, Partition, Table


class LDACollectiveMapper(CollectiveMapper):
    def setup(self, context: Context) -> None:
        c = context.get_configuration()
        self.K = int(c["num_topics"])
        self.alpha = float(c["alpha"])
        self.beta = float(c["beta"])
        self.num_iterations = int(c["num_iterations"])
        self.min_bound = int(c.get("min_bound", 100))
        self.max_bound = int(c.get("max_bound", 100))
        self.model_dir = c.get("model_dir")
        self.print_model = bool(c.get("print_model", False))
        self.print_interval = int(c.get("print_interval", 10))
        self.seed = int(c.get("seed", 0))
        self.blocks = c.get("blocks")

    def map_collective(self, reader: KeyValReader, context: Context) -> None:
        files = []
        while reader.next_key_value():
            files.append(reader.get_current_value())
        dev = self.device
        names, docs = F.read_lda_docs(files) if files else ([], [])
        self.doc_names = names
        if docs:
            d = torch.from_numpy(np.concatenate([np.full(len(x), i, dtype=np.int64) for i, x in enumerate(docs)])).to(dev)
            w = torch.from_numpy(np.concatenate(docs)).to(dev)
        else:
            d = torch.zeros(0, dtype=torch.long, device=dev); w = torch.zeros(0, dtype=torch.long, device=dev)
        t = Table(0, DoubleArrMax())
        t.add_partition(Partition(0, torch.tensor([float(w.max().item()) if w.numel() else -1.0], dtype=torch.float64, device=dev)))
        self.allreduce("lda", "allreduce-vocab", t)
        self.vocab = int(t.get_partition(0).get().item()) + 1
        self.engine = LDAEngine(d, w, len(docs), self.vocab, self.K, self.alpha, self.beta, symm=self.symm,
                                blocks=int(self.blocks) if self.blocks else None, seed=self.seed)
        eng = self.engine
        counts = self.comm.all_gather_object(int(d.numel()))
        total_tokens = max(sum(counts), 1)
        deadline_ms: Optional[float] = None
        bounded = self.min_bound < 100 or self.max_bound < 100
        self.likelihood_history: List[tuple] = []
        self.trained_history: List[int] = []
        for it in range(1, self.num_iterations + 1):
            t0 = time.perf_counter()
            n = eng.iterate(deadline_ms)
            eng.synchronize()
            dt = time.perf_counter() - t0
            n_all = sum(self.comm.all_gather_object(int(n.item())))
            self.trained_history.append(n_all)
            if bounded:
                pct = 100.0 * n_all / total_tokens
                if deadline_ms is None:
                    # first bounded guess from the measured full-pass rate
                    deadline_ms = 1e3 * dt / (2 * self.get_num_workers()) * self.max_bound / 100.0
                elif pct < self.min_bound:
                    deadline_ms *= 2.0
                elif pct > self.max_bound:
                    deadline_ms /= 2.0
            if it % self.print_interval == 0 or it == self.num_iterations:
                ll = eng.log_likelihood()
                self.likelihood_history.append((it, ll))
                if self.is_master():
                    LOG.info("iteration %d log-likelihood %.6e (%.3f s/iter)", it, ll, dt)
                if self.print_model and self.model_dir:
                    self.save_word_model(it)
            context.progress()
        self.log_likelihood = self.likelihood_history[-1][1] if self.likelihood_history else eng.log_likelihood()
        if self.model_dir and self.is_master():
            os.makedirs(self.model_dir, exist_ok=True)
            with open(os.path.join(self.model_dir, "evaluation"), "w") as f:
                f.write(f"{self.log_likelihood}\n")

    def save_word_model(self, it: int) -> None:
        ids, rows = self.engine.home_word_rows()
        d = os.path.join(self.model_dir, "tmp_word_model", str(it))
        os.makedirs(d, exist_ok=True)
        F.write_lda_word_model(os.path.join(d, str(self.get_self_id())), ids.cpu().tolist(), rows.cpu().numpy())
