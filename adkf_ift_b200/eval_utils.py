"""The step before and after the hot path, host side: task -> (support, query) sampling and the evaluation loop with its
result CSVs, with the reference's names (fs_mol/data/fsmol_task.py:100-144, fs_mol/data/fsmol_task_sampler.py:299-440,
fs_mol/utils/test_utils.py:165-350).  A "dataset" is anything with `get_task_reading_iterable(data_fold) -> Iterable[FSMolTask]`
(the reference's FSMolDataset has that method; `synthetic.SyntheticFSMolDataset` is the one used by the tests)."""
import csv
import dataclasses
import logging
import os
from dataclasses import dataclass
from enum import IntEnum
from typing import Callable, Dict, Iterable, Iterator, List, Optional, Tuple, Union

import numpy as np

from .fsmol_batcher import FSMolTaskSample, MoleculeDatapoint
from .metrics import BinaryEvalMetrics, NumericEvalMetrics

logger = logging.getLogger(__name__)


class DataFold(IntEnum):            # fs_mol/data/fsmol_dataset.py:21-24
    TRAIN = 0
    VALIDATION = 1
    TEST = 2


@dataclass
class FSMolTask:                    # fs_mol/data/fsmol_task.py:100-117
    name: str
    samples: List[MoleculeDatapoint]

    def get_pos_neg_separated(self) -> Tuple[List[MoleculeDatapoint], List[MoleculeDatapoint]]:
        """fs_mol/data/fsmol_task.py:96-98 unpacks `more_itertools.partition(pred, samples)` as (pos, neg), but partition
        yields the items with pred == False FIRST: the reference's "pos_samples" are the inactive molecules and its
        "neg_samples" the active ones.  The sampler below concatenates neg + pos and labels them 0.. / 1.., so the order of
        the molecules entering StratifiedShuffleSplit -- and therefore every support / query draw -- depends on this.  Kept as
        is (pinned by tests/golden/eval_harness.npz against the reference's own sampler)."""
        first = [s for s in self.samples if not s.bool_label]
        second = [s for s in self.samples if s.bool_label]
        return first, second

    @staticmethod
    def load_from_file(path) -> "FSMolTask":          # fs_mol/data/fsmol_task.py:108-144
        from .fsmol_dataset import load_task_from_file
        return load_task_from_file(str(path))


class SamplingException(Exception):
    def __init__(self, task_name: str, num_samples: int, num_train: int, num_valid: int = 0, num_test: Optional[int] = None):
        super().__init__(f"cannot sample {num_train} train / {num_valid} valid / {num_test} test from the {num_samples} "
                         f"datapoints of task {task_name}")
        self.task_name, self.num_samples, self.num_train, self.num_valid, self.num_test = task_name, num_samples, num_train, num_valid, num_test


class DatasetTooSmallException(SamplingException):
    pass


class DatasetClassTooSmallException(SamplingException):
    pass


class FoldTooSmallException(SamplingException):
    def __init__(self, task_name: str, num_samples: int, fold_name: str, num_train: int, num_valid: int = 0,
                 num_test: Optional[int] = None):
        super().__init__(task_name, num_samples, num_train, num_valid, num_test)
        self.fold_name = fold_name


class StratifiedTaskSampler:
    """fsmol_task_sampler.py:299-440.  Like the reference it defers the split itself to scikit-learn's
    StratifiedShuffleSplit with `random_state=seed`, so the same (task, seed) gives the same support / query molecules."""

    def __init__(self, train_size_or_ratio: Union[int, float] = 128, valid_size_or_ratio: Union[int, float] = 0.0,
                 test_size_or_ratio: Optional[Union[int, float, Tuple[int, int]]] = 256, allow_smaller_test: bool = True):
        self._train_size_or_ratio = train_size_or_ratio
        self._valid_size_or_ratio = valid_size_or_ratio
        self._test_size_or_ratio = test_size_or_ratio
        self._allow_smaller_test = allow_smaller_test

    def sample(self, task: FSMolTask, seed: int = 0) -> FSMolTaskSample:
        from sklearn.model_selection import StratifiedShuffleSplit
        pos, neg = task.get_pos_neg_separated()
        samples = neg + pos
        n = len(samples)
        labels = np.concatenate((np.zeros(len(neg)), np.ones(len(pos))))
        if isinstance(self._train_size_or_ratio, int):
            possible_test = n - self._train_size_or_ratio
        else:
            possible_test = n - int(n * self._train_size_or_ratio)
        t = self._test_size_or_ratio
        if t is None:
            num_test = possible_test
        else:
            if isinstance(t, int):
                num_test = t
            elif isinstance(t, tuple):
                num_test = max(t[0], min(t[1], possible_test))
            else:
                num_test = int(t * n)
            if self._allow_smaller_test:
                num_test = min(num_test, possible_test)
        if num_test < 2:
            raise DatasetTooSmallException(task.name, n, self._train_size_or_ratio, 0, num_test)
        split = StratifiedShuffleSplit(n_splits=1, train_size=self._train_size_or_ratio, test_size=num_test, random_state=seed)
        train_valid_idx, test_idx = next(iter(split.split(X=np.arange(n), y=labels)))
        train_valid = [samples[i] for i in train_valid_idx]
        test = [samples[i] for i in test_idx]
        if len(test) < 2:
            raise FoldTooSmallException(task.name, n, "test", len(train_valid))
        if self._valid_size_or_ratio > 0:
            split2 = StratifiedShuffleSplit(n_splits=1, test_size=self._valid_size_or_ratio, random_state=seed)
            tr, va = next(iter(split2.split(X=np.arange(len(train_valid)), y=[s.bool_label for s in train_valid])))
            train, valid = [train_valid[i] for i in tr], [train_valid[i] for i in va]
            if not 0 < sum(s.bool_label for s in valid) < len(valid):
                raise FoldTooSmallException(task.name, n, "valid", len(train), num_test=len(test))
        else:
            train, valid = train_valid, []
        if not 0 < sum(s.bool_label for s in train) < len(train):
            raise FoldTooSmallException(task.name, n, "train", len(train), num_test=len(test))
        if not 0 < sum(s.bool_label for s in test) < len(test):
            raise FoldTooSmallException(task.name, n, "test", len(train), num_test=len(test))
        return FSMolTaskSample(name=task.name, train_samples=train, valid_samples=valid, test_samples=test)


@dataclass(frozen=True)
class FSMolTaskSampleEvalResults(BinaryEvalMetrics):              # test_utils.py:42-49
    task_name: str = ""
    seed: int = 0
    num_train: int = 0
    num_test: int = 0
    fraction_pos_train: float = 0.0
    fraction_pos_test: float = 0.0


@dataclass(frozen=True)
class FSMolTaskSampleEvalResultsNumeric(NumericEvalMetrics):      # test_utils.py:52-57
    task_name: str = ""
    seed: int = 0
    num_train: int = 0
    num_test: int = 0


def write_csv_summary(output_csv_file: str, test_results: Iterable[FSMolTaskSampleEvalResults]):
    """Column names and order of test_utils.py:165-207 (what collect_eval_runs.py reads)."""
    fieldnames = ["num_train_requested", "num_train", "fraction_positive_train", "num_test", "fraction_positive_test", "seed",
                  "valid_score", "average_precision_score", "roc_auc", "acc", "balanced_acc", "precision", "recall", "f1_score",
                  "delta_auprc"]
    with open(output_csv_file, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=fieldnames)
        w.writeheader()
        for r in test_results:
            w.writerow({"num_train_requested": r.num_train, "num_train": r.num_train, "num_test": r.num_test,
                        "fraction_positive_train": r.fraction_pos_train, "fraction_positive_test": r.fraction_pos_test,
                        "seed": r.seed, "average_precision_score": r.avg_precision, "roc_auc": r.roc_auc, "acc": r.acc,
                        "balanced_acc": r.balanced_acc, "precision": r.prec, "recall": r.recall, "f1_score": r.f1,
                        "delta_auprc": r.avg_precision - r.fraction_pos_test})


def write_csv_summary_numeric(output_csv_file: str, test_results: Iterable[FSMolTaskSampleEvalResultsNumeric]):
    fieldnames = ["num_train_requested", "num_train", "num_test", "seed", "mse", "mae", "r2"]     # test_utils.py:210-233
    with open(output_csv_file, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=fieldnames)
        w.writeheader()
        for r in test_results:
            w.writerow({"num_train_requested": r.num_train, "num_train": r.num_train, "num_test": r.num_test, "seed": r.seed,
                        "mse": r.mse, "mae": r.mae, "r2": r.r2})


def numeric_task_is_filtered(task: FSMolTask) -> bool:
    """`filter_numeric_labels` rule of test_utils.py:281-290 / dkt.py:172-181: percentage-like, non-positive, inf or nan
    measurements are not regression tasks (the labels go through a log)."""
    y = np.array([s.numeric_label for s in task.samples], dtype=np.float64)
    return bool((np.all(y >= 0.0) and np.all(y <= 100.0)) or np.any(y <= 0.0) or np.any(np.isinf(y)) or np.any(np.isnan(y)))


@dataclass
class EvalDraw:
    """One (task, support size, run) of the evaluation protocol."""
    task_name: str
    train_size: int
    seed: int
    sample: FSMolTaskSample


def iter_eval_draws(dataset, train_set_sample_sizes: List[int], num_samples: int = 10,
                    valid_size_or_ratio: Union[int, float] = 0.0,
                    test_size_or_ratio: Optional[Union[int, float, Tuple[int, int]]] = None, fold: DataFold = DataFold.TEST,
                    seed: int = 0, filter_numeric_labels: bool = False) -> Iterator[Tuple[FSMolTask, List[EvalDraw]]]:
    """The sampling half of eval_model (test_utils.py:277-318): per task, every (support size, run) draw in the reference's
    order with `local_seed = seed + run_idx`; draws that cannot be made are skipped, filtered tasks are not yielded."""
    for task in dataset.get_task_reading_iterable(fold):
        if filter_numeric_labels and numeric_task_is_filtered(task):
            continue
        draws: List[EvalDraw] = []
        for train_size in train_set_sample_sizes:
            sampler = StratifiedTaskSampler(train_size_or_ratio=train_size, valid_size_or_ratio=valid_size_or_ratio,
                                            test_size_or_ratio=test_size_or_ratio, allow_smaller_test=True)
            for run_idx in range(num_samples):
                local_seed = seed + run_idx
                try:
                    draws.append(EvalDraw(task.name, train_size, local_seed, sampler.sample(task, seed=local_seed)))
                except (SamplingException, ValueError) as e:
                    logger.debug(f"Failed to draw sample with {train_size} train points for {task.name}. Skipping. {e}")
        yield task, draws


def _pos_ratio(samples: List[MoleculeDatapoint]) -> float:
    return float(np.mean([bool(s.bool_label) for s in samples])) if samples else 0.0


def make_eval_result(draw: EvalDraw, metrics, numeric: bool):
    if numeric:
        return FSMolTaskSampleEvalResultsNumeric(task_name=draw.task_name, seed=draw.seed, num_train=draw.train_size,
                                                 num_test=len(draw.sample.test_samples), **dataclasses.asdict(metrics))
    return FSMolTaskSampleEvalResults(task_name=draw.task_name, seed=draw.seed, num_train=draw.train_size,
                                      num_test=len(draw.sample.test_samples),
                                      fraction_pos_train=_pos_ratio(draw.sample.train_samples),
                                      fraction_pos_test=_pos_ratio(draw.sample.test_samples), **dataclasses.asdict(metrics))


def write_task_results(out_dir: Optional[str], task_name: str, results: List, numeric: bool) -> None:
    if out_dir is None:
        return
    os.makedirs(out_dir, exist_ok=True)
    path = os.path.join(out_dir, f"{task_name}_eval_results.csv")
    (write_csv_summary_numeric if numeric else write_csv_summary)(path, results)


def eval_model(test_model_fn: Callable[[FSMolTaskSample, str, int], object], dataset, train_set_sample_sizes: List[int],
               out_dir: Optional[str] = None, num_samples: int = 10, valid_size_or_ratio: Union[int, float] = 0.0,
               test_size_or_ratio: Optional[Union[int, float, Tuple[int, int]]] = None, fold: DataFold = DataFold.TEST,
               task_reader_fn=None, seed: int = 0, filter_numeric_labels: bool = False) -> Dict[str, List]:
    """test_utils.py:236-350 with a per-sample callback (the reference's calling convention; `evaluate_adkt_model` batches
    the draws of many tasks into single launches instead)."""
    import tempfile
    if task_reader_fn is not None:
        raise NotImplementedError("task_reader_fn: pass a dataset whose get_task_reading_iterable applies the transformation")
    task_to_results: Dict[str, List] = {}
    for task, draws in iter_eval_draws(dataset, train_set_sample_sizes, num_samples, valid_size_or_ratio, test_size_or_ratio, fold,
                                       seed, filter_numeric_labels):
        results = []
        for d in draws:
            with tempfile.TemporaryDirectory() as tmp:
                results.append(make_eval_result(d, test_model_fn(d.sample, tmp, d.seed), filter_numeric_labels))
        task_to_results[task.name] = results
        write_task_results(out_dir, task.name, results, filter_numeric_labels)
    return task_to_results
