"""``Initialize`` / ``HarpDAALConstants`` / ``Service`` / ``DataGenerator``
(``core/harp-daal-interface/src/main/java/edu/iu/{data_aux,data_gen}``).

``load_sys_args`` keeps the six standard positional arguments of every Harp-DAAL launcher —
``<num_mappers> <threads> <mem_MB> <iterations> <inputDir> <workDir>`` (``Initialize.java:97-129``).
``loadDistributedLibs`` (shipping DAAL/TBB/OMP ``.so`` files through the DistributedCache,
``:67-83``) has no work to do here: the native library travels with the package.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Any, Dict, List, Sequence

import numpy as np
import torch

from ..io import datagen, formats as F


class HarpDAALConstants:
    """Configuration keys (``HarpDAALConstants.java:19-71``)."""
    NUM_MAPPERS = "num_mappers"
    NUM_THREADS = "num_threads"
    NUM_ITERATIONS = "num_iterations"
    WORK_DIR = "work_dir"
    FILE_DIM = "file_dim"
    FEATURE_DIM = "feature_dim"
    NUM_CENTROIDS = "num_centroids"
    NUM_CLASS = "num_class"
    NUM_DEPVAR = "num_depvar"
    NUM_FACTOR = "num_factor"
    BATCH_SIZE = "batch_size"
    LEARNING_RATE = "learning_rate"
    TEST_FILE_PATH = "test_file_path"
    TEST_TRUTH_PATH = "test_truth_path"
    DO_GEN_DATA = "do_gen_data"


@dataclass
class SysArgs:
    num_mappers: int
    num_threads: int
    mem_mb: int
    num_iterations: int
    input_dir: str
    work_dir: str


class Initialize:
    SYS_ARG_NUM = 6

    def __init__(self, conf: Dict[str, Any], args: Sequence[str]):
        self.conf = conf
        self.args = list(args)
        self.sys: SysArgs = None

    def load_distributed_libs(self) -> None:
        """Nothing to distribute: ``libharp_b200.so`` is in-tree."""

    def load_sys_args(self) -> SysArgs:
        a = self.args
        if len(a) < self.SYS_ARG_NUM:
            raise SystemExit("Usage: <num_mappers> <threads> <mem_MB> <iterations> <inputDir> <workDir> [app args...]")
        self.sys = SysArgs(int(a[0]), int(a[1]), int(a[2]), int(a[3]), a[4], a[5])
        self.conf.update({HarpDAALConstants.NUM_MAPPERS: self.sys.num_mappers,
                          HarpDAALConstants.NUM_THREADS: self.sys.num_threads,
                          HarpDAALConstants.NUM_ITERATIONS: self.sys.num_iterations,
                          HarpDAALConstants.WORK_DIR: self.sys.work_dir})
        return self.sys

    def get_sys_arg_num(self) -> int:
        return self.SYS_ARG_NUM

    def app_args(self) -> List[str]:
        return self.args[self.SYS_ARG_NUM:]

    def create_job(self, job_name: str, launcher_cls, mapper_cls) -> "Job":
        """``Initialize.createJob`` (``Initialize.java:132-162``): a job that deals the files of ``inputDir`` to the
        workers like ``MultiFileInputFormat`` and runs ``mapper_cls`` on each.  The Hadoop knobs it sets there
        (framework name, memory, JVM options, task timeout, zero reducers) have no counterpart under torchrun; the
        number of map tasks is checked against the world size when the job runs."""
        if self.sys is None:
            self.load_sys_args()
        return Job(job_name, launcher_cls, mapper_cls, dict(self.conf), self.sys.input_dir,
                   os.path.join(self.sys.work_dir, "output"), self.sys.num_mappers)

    loadDistributedLibs, loadSysArgs, getSysArgNum, createJob = (load_distributed_libs, load_sys_args, get_sys_arg_num,
                                                                 create_job)


class Job:
    """What ``createJob`` returns: enough of ``org.apache.hadoop.mapreduce.Job`` for the reference's launchers
    (``job.waitForCompletion(true)``, ``job.isSuccessful()``)."""

    def __init__(self, name: str, launcher_cls, mapper_cls, conf: Dict[str, Any], input_path: str, output_path: str,
                 num_map_tasks: int):
        self.name, self.launcher_cls, self.mapper_cls = name, launcher_cls, mapper_cls
        self.conf, self.input_path, self.output_path, self.num_map_tasks = conf, input_path, output_path, num_map_tasks
        self.mapper = None
        self._ok = False

    def get_configuration(self) -> Dict[str, Any]:
        return self.conf

    def wait_for_completion(self, verbose: bool = True) -> bool:
        from ..launch.common import check_world, run_mapper
        check_world(self.num_map_tasks)
        self.mapper = self.mapper_cls()
        try:
            run_mapper(self.mapper, self.input_path, self.conf, self.output_path)
            self._ok = True
        except Exception:  # noqa: BLE001
            self._ok = False
            if verbose:
                raise
        return self._ok

    def is_successful(self) -> bool:
        return self._ok

    getConfiguration, waitForCompletion, isSuccessful = get_configuration, wait_for_completion, is_successful


class Service:
    """Table printers (``Service.java:69-659``)."""

    @staticmethod
    def print_numeric_table(header: str, t: torch.Tensor, max_rows: int = 10, max_cols: int = 10) -> str:
        t2 = t if t.dim() == 2 else t.view(1, -1)
        lines = [header]
        for row in t2[:max_rows]:
            lines.append(" ".join(f"{float(v):<12.6g}" for v in row[:max_cols]))
        out = "\n".join(lines)
        print(out)
        return out

    printNumericTable = print_numeric_table


class DataGenerator:
    """``DataGenerator.generateDenseDataMulti / Single`` (``DataGenerator.java:56-249``)."""

    @staticmethod
    def generate_dense_data_multi(num_rows: int, dim: int, num_files: int, out_dir: str, sep: str = ",", seed: int = 0) -> List[str]:
        return datagen.generate_dense_csv(num_rows, dim, num_files, out_dir, seed=seed)

    @staticmethod
    def generate_dense_data_single(num_rows: int, dim: int, path: str, sep: str = ",", seed: int = 0) -> str:
        rng = np.random.default_rng(seed)
        F.write_dense_csv(path, rng.standard_normal((num_rows, dim)), sep)
        return path

    generateDenseDataMulti, generateDenseDataSingle = generate_dense_data_multi, generate_dense_data_single
