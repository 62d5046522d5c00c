"""Worker: the select-device ring test on CPU buffers over gloo (rank r must receive r-1)."""
import os

from rocm_mpi_b200.models.selectdevice import selectdevice_test

me, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
out = selectdevice_test(4, quiet=True)
assert out["gloo"].tolist() == [float((me - 1) % world)] * 4
print(f"rank {me} ok")
