# facts about the GPU box that the host-side design depends on
echo "== cpu"; nproc; lscpu | grep -E "Model name|Socket|NUMA|Thread|Core" ; 
echo "== mem"; free -g | head -2; ulimit -l
echo "== cgroup"; cat /sys/fs/cgroup/cpu.max 2>/dev/null; cat /sys/fs/cgroup/memory.max 2>/dev/null
echo "== gpus"; nvidia-smi -L; nvidia-smi topo -m 2>/dev/null | head -30
echo "== nccl libs"; ls /usr/lib/x86_64-linux-gnu/ | grep -i nccl; python -c "import nvidia.nccl,os;print(os.listdir(os.path.dirname(nvidia.nccl.__file__)+'/lib'))" 2>/dev/null
