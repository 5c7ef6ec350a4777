PDSH_LAUNCHER = "pdsh"
PDSH_MAX_FAN_OUT = 1024
OPENMPI_LAUNCHER = "openmpi"
MVAPICH_LAUNCHER = "mvapich"
MVAPICH_TMP_HOSTFILE = "/tmp/deepspeed_mvapich_hostfile"
SLURM_LAUNCHER = "slurm"
MOSAICML_LAUNCHER = "mosaicml"
