"""xacc-vqe_b200: B200-native backend for the Pauli-Hamiltonian x statevector path of xacc-vqe.

Import as ``xaccvqe_b200`` (the repo-root shim maps that name onto this directory, whose name is
not a valid Python identifier)."""
from . import capi  # noqa: F401
from .capi import B200Error, Backend, OpHandle, uccsd_gates, uccsd_rotations  # noqa: F401
from . import tasks  # noqa: F401
from .tasks import (AcceleratorBuffer, Ansatz, ComputeEnergyVQETask, DiagonalizeTask, GateAnsatz, PauliOperator,  # noqa: F401
                    RDMGenerator, VQEMinimizeTask, VQEProgram, VQETaskResult, compile, execute, nelder_mead, rdm_energy)
