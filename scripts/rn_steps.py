"""Run a few device-resident ResNet-50 sync-PS steps on cuda:0 and nothing else, for ncu launch
lists (set TFGRAFT_GRAPH=0 so each kernel is a launch of its own):
    ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv \
        python scripts/rn_steps.py 256 3
The per-launch durations are then matched to layers by launch order (scripts/rn_layers.py)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch

    import bench
    from k8s_b200.syncps import SyncPSJob, synthetic_images

    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    image = int(sys.argv[3]) if len(sys.argv) > 3 else 224
    torch.cuda.set_device(0)
    stream = torch.cuda.Stream()
    job = SyncPSJob(bench.tf_config_for_rank(1, 0), 0, 1, cuda_device=0, stream=stream.cuda_stream)
    job.build_resnet50(batch, image, 1000)
    job.optimizer("momentum", 0.001, 0.9)
    job.connect_peers(None)
    x, y = synthetic_images(0, batch, image, 1000, 0)
    dx, dy = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    for _ in range(steps):
        job.step_dev([dx], [dy])
    job.sync()
    print("loss", job.loss(0), "launches", job.launches)
    job.close()


if __name__ == "__main__":
    main()
