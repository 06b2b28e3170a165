"""Drop-in for KubeRay's `fine-tune-pytorch-text-classifier.py` on the b200flow runtime.

The reference submits the upstream script as a RayJob
(examples/ray/lightning-bert/fine-tune.yaml:46-49).  Same recipe ([UP]):
GLUE-CoLA-shaped data, `bert-base-cased`-shaped classifier, a Lightning-style module
(training_step / validation_step / configure_optimizers), DDP across the worker group,
Matthews correlation on validation each epoch, `train.report(metrics, checkpoint)` and
`CheckpointConfig(num_to_keep=2, checkpoint_score_attribute="matthews_correlation")`.
Ray's and Lightning's APIs are provided by `b200flow.engines.ray_train_api` /
`lightning_lite`; data is synthetic (no network), weights random-init.

Size knobs: B2F_MODEL, B2F_EPOCHS, B2F_TRAIN_SAMPLES, B2F_EVAL_SAMPLES.
"""
import os

import torch

from b200flow.data.synthetic import cola_like, collate_pad
from b200flow.engines import ray_train_api as train
from b200flow.engines.lightning_lite import (LightningModule, RayDDPStrategy, RayLightningEnvironment,
                                             RayTrainReportCallback, Trainer, prepare_trainer)
from b200flow.engines.ray_train_api import (CheckpointConfig, ListDataset, RunConfig, ScalingConfig,
                                            TorchTrainer)
from b200flow.models.bert import BertForSequenceClassification, bert_config
from b200flow.optim import AdamW


def matthews_corrcoef(preds, refs) -> float:
    p, r = torch.as_tensor(preds).long(), torch.as_tensor(refs).long()
    tp = float(((p == 1) & (r == 1)).sum()); tn = float(((p == 0) & (r == 0)).sum())
    fp = float(((p == 1) & (r == 0)).sum()); fn = float(((p == 0) & (r == 1)).sum())
    den = ((tp + fp) * (tp + fn) * (tn + fp) * (tn + fn)) ** 0.5
    return (tp * tn - fp * fn) / den if den > 0 else 0.0


class SentimentModel(LightningModule):
    def __init__(self, lr=2e-5, eps=1e-8):
        super().__init__()
        self.lr, self.eps = lr, eps
        self.cfg = bert_config(os.environ.get("B2F_MODEL", "bert-base"), num_labels=2)
        self.model = BertForSequenceClassification(self.cfg, dtype=torch.float32)
        self.predictions, self.references = [], []

    def forward(self, batch):
        return self.model(batch["input_ids"], batch.get("token_type_ids"), batch["attention_mask"],
                          batch.get("labels"))

    def training_step(self, batch, batch_idx):
        loss = self.forward(batch)["loss"]
        self.log("train_loss", loss)
        return loss

    def validation_step(self, batch, batch_idx):
        out = self.forward(batch)
        self.predictions.append(out["logits"].argmax(-1).cpu())
        self.references.append(batch["labels"].cpu())

    def on_validation_epoch_end(self):
        mcc = matthews_corrcoef(torch.cat(self.predictions), torch.cat(self.references))
        self.log("matthews_correlation", mcc, sync_dist=True)
        self.predictions.clear()
        self.references.clear()

    def configure_optimizers(self):
        return AdamW(self.parameters(), lr=self.lr, eps=self.eps)


def train_func(config):
    lr, eps, batch_size, max_epochs = config["lr"], config["eps"], config["batch_size"], config["max_epochs"]
    train_ds = train.get_dataset_shard("train")
    val_ds = train.get_dataset_shard("validation")
    train_dl = lambda: train_ds.iter_torch_batches(batch_size=batch_size, collate_fn=collate_pad)  # noqa: E731
    val_dl = lambda: val_ds.iter_torch_batches(batch_size=batch_size, collate_fn=collate_pad)      # noqa: E731
    model = SentimentModel(lr=lr, eps=eps)
    trainer = Trainer(max_epochs=max_epochs, accelerator="auto", devices="auto", strategy=RayDDPStrategy(),
                      plugins=[RayLightningEnvironment()], callbacks=[RayTrainReportCallback()],
                      enable_progress_bar=False)
    trainer = prepare_trainer(trainer)
    trainer.fit(model, train_dataloaders=train_dl, val_dataloaders=val_dl)


if __name__ == "__main__":
    nt = int(os.environ.get("B2F_TRAIN_SAMPLES", 0)) or None
    nv = int(os.environ.get("B2F_EVAL_SAMPLES", 0)) or None
    vocab = bert_config(os.environ.get("B2F_MODEL", "bert-base")).vocab_size
    ray_datasets = {"train": ListDataset.from_items(cola_like("train", vocab=vocab)[:nt]),
                    "validation": ListDataset.from_items(cola_like("validation", vocab=vocab)[:nv])}
    use_gpu = torch.cuda.is_available()
    num_workers = int(os.environ.get("B2F_RAY_WORKERS", 1))
    run_config = RunConfig(
        name="ptl-sent-classification", storage_path=os.path.join(os.environ.get("HOME", "/tmp"), "ray_results"),
        checkpoint_config=CheckpointConfig(num_to_keep=2, checkpoint_score_attribute="matthews_correlation",
                                           checkpoint_score_order="max"))
    trainer = TorchTrainer(
        train_func,
        train_loop_config={"lr": 1e-5, "eps": 1e-8, "batch_size": 16,
                           "max_epochs": int(os.environ.get("B2F_EPOCHS", 5))},
        scaling_config=ScalingConfig(num_workers=num_workers, use_gpu=use_gpu),
        run_config=run_config, datasets=ray_datasets)
    result = trainer.fit()
    print("Final metrics:", result.metrics)
    print("Best checkpoints:", [(c.path, m.get("matthews_correlation")) for c, m in result.best_checkpoints])
