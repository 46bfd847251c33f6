"""What is being computed: the filter table and the three reference programs as presets."""
from .filters import FILTERS, Filter
from .programs import PROGRAMS, CudaProgram, MpiProgram, OmpProgram, Program

__all__ = ["FILTERS", "Filter", "PROGRAMS", "Program", "MpiProgram", "OmpProgram", "CudaProgram"]
