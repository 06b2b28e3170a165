"""Lightning-shaped trainer (b200flow/engines/lightning_lite.py): optimizer / scheduler return shapes, accumulation,
clipping and step limits, single process."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from b200flow.engines.lightning_lite import Callback, LightningModule, Trainer, _unpack_optimizers  # noqa: E402


class Toy(LightningModule):
    def __init__(self, shape="dict"):
        super().__init__()
        torch.manual_seed(0)
        self.net = torch.nn.Linear(4, 1)
        self.shape, self.epochs_ended, self.val_seen = shape, 0, 0

    def training_step(self, batch, i):
        loss = torch.nn.functional.mse_loss(self.net(batch["x"]), batch["y"])
        self.log("train_loss", loss)
        return loss

    def validation_step(self, batch, i):
        self.val_seen += 1

    def on_train_epoch_end(self):
        self.epochs_ended += 1

    def configure_optimizers(self):
        opt = torch.optim.SGD(self.parameters(), lr=0.1)
        sch = torch.optim.lr_scheduler.StepLR(opt, step_size=1, gamma=0.5)
        return {"dict": {"optimizer": opt, "lr_scheduler": {"scheduler": sch, "interval": "step"}},
                "tuple": ([opt], [sch]), "plain": opt, "list": [opt]}[self.shape]


def _data(n=8):
    g = torch.Generator().manual_seed(1)
    return [{"x": torch.randn(4, 4, generator=g), "y": torch.randn(4, 1, generator=g)} for _ in range(n)]


def test_configure_optimizers_return_shapes():
    for shape, has_sched, interval in (("dict", True, "step"), ("tuple", True, "epoch"), ("plain", False, "epoch"),
                                       ("list", False, "epoch")):
        opt, sch, iv = _unpack_optimizers(Toy(shape).configure_optimizers())
        assert isinstance(opt, torch.optim.SGD) and (sch is not None) == has_sched and iv == interval


def test_accumulation_clipping_limits_and_hooks():
    seen = []

    class Spy(Callback):
        def on_train_epoch_end(self, trainer, module):
            seen.append((trainer.current_epoch, trainer.global_step, module._logged["train_loss"]))

    data = _data(8)
    m = Toy("dict")
    t = Trainer(max_epochs=2, accumulate_grad_batches=2, gradient_clip_val=0.5, limit_val_batches=3, callbacks=[Spy()])
    t.fit(m, train_dataloaders=lambda: iter(data), val_dataloaders=lambda: iter(data))
    assert t.global_step == 8 and m.epochs_ended == 2 and m.val_seen == 6
    assert [s[:2] for s in seen] == [(0, 4), (1, 8)] and seen[1][2] < seen[0][2]
    # accumulate 2 == one batch of the two micro-batches concatenated
    a, b = Toy("plain"), Toy("plain")
    Trainer(max_epochs=1, accumulate_grad_batches=2).fit(a, train_dataloaders=data[:4])
    big = [{k: torch.cat([data[i][k], data[i + 1][k]]) for k in ("x", "y")} for i in (0, 2)]
    Trainer(max_epochs=1).fit(b, train_dataloaders=big)
    assert torch.allclose(a.net.weight, b.net.weight, atol=1e-6)
    # max_steps and limit_train_batches cut the loop; an odd tail window still steps
    c = Toy("plain")
    tc = Trainer(max_epochs=5, max_steps=3)
    tc.fit(c, train_dataloaders=data)
    assert tc.global_step == 3 and c.epochs_ended == 1
    d = Toy("plain")
    td = Trainer(max_epochs=1, accumulate_grad_batches=2, limit_train_batches=3)
    td.fit(d, train_dataloaders=data)
    assert td.global_step == 2
