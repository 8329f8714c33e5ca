"""Launcher constants (reference: deepspeed/launcher/constants.py)."""
PDSH_LAUNCHER = "pdsh"
PDSH_MAX_FAN_OUT = 1024
OPENMPI_LAUNCHER = "openmpi"
MPICH_LAUNCHER = "mpich"
IMPI_LAUNCHER = "impi"
SLURM_LAUNCHER = "slurm"
MVAPICH_LAUNCHER = "mvapich"
MVAPICH_TMP_HOSTFILE = "/tmp/sxb200_mvapich_hostfile"
ELASTIC_TRAINING_ID_DEFAULT = "123456789"
PID_FILE_BASEPATH = "/tmp"
TORCH_DISTRIBUTED_DEFAULT_PORT = 29500
