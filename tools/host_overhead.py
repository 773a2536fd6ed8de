"""Host-side cost of one kfb_factor_update / kfb_precondition call on the ResNet-50 job set
(time until the call returns, nothing synchronised) next to its device time."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import bench  # noqa: E402


def main():
    args = bench.parse_args()
    device = torch.device('cuda:0')
    torch.cuda.set_device(device)
    model, pre, x, y = bench.build(args, device) if hasattr(bench, 'build') else (None, None, None, None)
    print('bench.build missing' if pre is None else 'ok')


if __name__ == '__main__':
    main()
