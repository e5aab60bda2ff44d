"""CPU runs: the intra-op thread count follows the model's layer sizes (``elephas_b200/utils/cpu_threads.py``)."""
import torch

from elephas_b200.models import SGD, zoo
from elephas_b200.utils import cpu_threads


def test_small_layers_run_on_one_thread_and_the_choice_is_adjustable(monkeypatch):
    start = torch.get_num_threads()
    try:
        model = zoo.mnist_mlp()                      # 784-128-128-10: a step's largest product is 13 MFLOP at batch 64
        model.compile(SGD(learning_rate=0.1), 'categorical_crossentropy', ['acc'])
        assert cpu_threads.tune_for(model, 64) == 1 and torch.get_num_threads() == 1
        monkeypatch.setenv('ELEPHAS_B200_CPU_THREADS', '2')
        assert cpu_threads.tune_for(model, 64) == 2 and torch.get_num_threads() == 2
        monkeypatch.setenv('ELEPHAS_B200_CPU_THREADS', 'default')
        torch.set_num_threads(start)
        assert cpu_threads.tune_for(model, 64) == start and torch.get_num_threads() == start
        monkeypatch.setenv('ELEPHAS_B200_CPU_THREADS', 'auto')
        wide = zoo.wide_mlp(input_dim=512, hidden=512, depth=2, out_dim=1)   # 2 * 4096 * 512 * 512 = 2 GFLOP: timed
        wide.compile(SGD(learning_rate=0.01), 'mse')
        chosen = cpu_threads.tune_for(wide, 4096)
        upper = cpu_threads.tune_for._upper   # the count torch had when this process first asked (other tests may have run)
        assert 1 <= chosen <= upper and torch.get_num_threads() == chosen
        assert cpu_threads.tune_for(wide, 4096) == chosen                    # cached per shape
    finally:
        torch.set_num_threads(start)
