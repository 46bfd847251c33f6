"""The reference's three programs as presets of the one engine.

==============  =======================  ==========================================================
preset          reference program        what it fixes
==============  =======================  ==========================================================
``mpi``         mpi/mpi_convolution.c    zero boundary, one rank per GPU on the RowsDivision grid,
                                         prints the max-over-ranks loop seconds as ``"%f\\n"``
``omp``         open-mp/omp_convolution.c same decomposition; the intra-rank thread level (the two
                                         ``#pragma omp parallel for``, :292,297) is the GPU's
                                         SIMT threads — there is nothing else to configure
``cuda``        cuda/                    single GPU; ``compat=True`` reproduces the frozen-border /
                                         odd-repetition behaviour of cuda_convolution.cu, default is
                                         the well-defined zero boundary; prints
                                         ``"Execution time: %.3f sec"`` (cuda/main.c:49)
==============  =======================  ==========================================================
"""
from __future__ import annotations

from dataclasses import dataclass, field


@dataclass
class Program:
    name: str
    boundary: str = "zero"
    gpus: int = 0                 # 0 = every visible GPU (the `mpiexec -n` analogue)
    report: str = "mpi"           # which reference stdout line to print
    extra: dict = field(default_factory=dict)

    def launcher_args(self, image, width, height, repetitions, color):
        a = [str(image), str(width), str(height), str(repetitions), color, "--boundary", self.boundary]
        if self.gpus:
            a += ["--gpus", str(self.gpus)]
        for k, v in self.extra.items():
            a += [f"--{k}", str(v)]
        return a

    def run(self, image, width, height, repetitions, color, cwd=None):
        """Run through the native launcher; returns its CompletedProcess."""
        from ..utils.cli import run_launcher

        return run_launcher(self.launcher_args(image, width, height, repetitions, color), cwd=cwd)


def MpiProgram(gpus=0, **extra):
    return Program("mpi", "zero", gpus, "mpi", extra)


def OmpProgram(gpus=0, **extra):
    return Program("omp", "zero", gpus, "mpi", extra)


def CudaProgram(compat=False, **extra):
    return Program("cuda", "cuda-compat" if compat else "zero", 1, "cuda", extra)


PROGRAMS = {"mpi": MpiProgram, "omp": OmpProgram, "cuda": CudaProgram}
