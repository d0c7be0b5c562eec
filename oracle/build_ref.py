"""oracle/build_ref.py -- ORACLE / TEST INFRASTRUCTURE ONLY.

Stages the reference's OWN Python files for the hot path, UNMODIFIED, under oracle/_ref/ (git-ignored, NOT gpurun-ignored:
it travels to the GPU box like a built .so; nothing under it enters the repository's history) so that
`bench.py --impl reference` and the `cpu_baseline` leg can time the reference's own `ADKTModel`, `run_on_batches`,
`functional_call` and `cauchy_hypergradient` on the GPU box's host cores -- `cpu_baseline.kind = "reference"` -- instead of
the oracle's restatement.  The reference is pure Python with no build system (no setup.py / pyproject build target), so there
is nothing to compile and `pip install --target` has nothing to install: this copy IS its build.  Third-party libraries the
files import (gpytorch, botorch, torch_scatter, ...) still come from oracle/shims.py (restated) unless really installed.

    python -m oracle.build_ref            (authoring container: needs /root/reference)

Called by __graft_entry__.build() when /root/reference exists; on the GPU box only the staged copy is used.
"""
import os
import shutil

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
SOURCE = os.environ.get("ADKF_REFERENCE_ROOT", "/root/reference")

# the files SURVEY.md section 8(a) cites (rows a1-a25), plus the package files they import
TREES = [
    ("fs_mol", ("__init__.py",)),
    ("fs_mol/models", ("__init__.py", "adaptive_dkt.py", "abstract_torch_fsmol_model.py", "dkt.py", "dkl.py")),
    ("fs_mol/modules", None),
    ("fs_mol/utils", ("__init__.py", "adaptive_dkt_utils.py", "cauchy_hypergradient.py", "cauchy_hypergradient_jvp.py",
                      "_stateless.py", "gp_utils.py", "torch_utils.py", "metrics.py", "metric_logger.py", "test_utils.py",
                      "logging.py", "cli_utils.py", "dkt_utils.py", "dkl_utils.py")),
    ("fs_mol/data", None),
    ("MoleculeNet/chem_lib", None),
    ("MoleculeNet/chem_lib/models", None),
    ("bayes_opt", ("bo_utils.py",)),
]


def build(force: bool = False) -> str:
    if not os.path.isdir(os.path.join(SOURCE, "fs_mol")):
        if os.path.isdir(os.path.join(DEST, "fs_mol")):
            return DEST                                   # GPU box: use what travelled
        raise RuntimeError(f"reference not found at {SOURCE} and nothing staged under {DEST}")
    stamp = os.path.join(DEST, ".staged_from")
    if not force and os.path.exists(stamp):
        return DEST
    if os.path.isdir(DEST):
        shutil.rmtree(DEST)
    n = 0
    for rel, names in TREES:
        src_dir = os.path.join(SOURCE, rel)
        if not os.path.isdir(src_dir):
            continue
        os.makedirs(os.path.join(DEST, rel), exist_ok=True)
        for name in (names if names is not None else sorted(os.listdir(src_dir))):
            src = os.path.join(src_dir, name)
            if os.path.isfile(src) and name.endswith(".py"):
                shutil.copy2(src, os.path.join(DEST, rel, name))
                n += 1
    with open(stamp, "w") as f:
        f.write(f"{SOURCE}\n{n} files, unmodified\n")
    return DEST


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv))
