# scratch: the command list of the last ad-hoc gpurun call (see tools/gpu_job.sh for the jobs of a round)
