#!/usr/bin/env python
"""Multi-worker synchronous training of the MNIST CNN - what the reference repository's single script does
(/root/reference/tf_dist_example.py: cluster from TF_CONFIG, MultiWorkerMirroredStrategy(AUTO), tf.data pipeline
batched by the global batch size with AutoShardPolicy.OFF, Sequential CNN built under strategy.scope(), fit for
10 epochs x 20 steps), written against this framework's API.

Unlike the reference, the cluster is NOT hard-coded into the script: TF_CONFIG comes from the environment
(README.md:158-162 of the reference) or from --tf-config / --workers + --index. Examples:

  # two workers on one machine, CPU, TCP ring all-reduce
  python examples/launch_local_cluster.py --workers 2 -- python examples/mnist_multiworker.py --device cpu
  # one process per GPU of this host, fused NVLink all-reduce
  python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 examples/mnist_multiworker.py --mirrored
  # a real cluster: on node i
  TF_CONFIG='{"cluster": {"worker": ["10.0.0.5:12345", "10.0.0.6:12345"]}, "task": {"type": "worker", "index": i}}' \\
      python examples/mnist_multiworker.py
"""
from __future__ import annotations

import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import tfdl  # noqa: E402

PER_WORKER_BATCH = 64
SHUFFLE_BUFFER = 10000


def parse_args() -> argparse.Namespace:
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("--tf-config", help="TF_CONFIG JSON (default: the environment variable)")
    ap.add_argument("--workers", help="comma-separated host:port list; with --index builds TF_CONFIG")
    ap.add_argument("--index", type=int, default=0)
    ap.add_argument("--mirrored", action="store_true", help="MirroredStrategy (one host; under torchrun: one replica per process)")
    ap.add_argument("--communication", default="AUTO", choices=["AUTO", "RING", "NCCL"])
    ap.add_argument("--device", default=None, help="cpu | gpu:K (default: every visible GPU, else CPU)")
    ap.add_argument("--epochs", type=int, default=10)
    ap.add_argument("--steps-per-epoch", type=int, default=20)
    ap.add_argument("--learning-rate", type=float, default=0.001)
    ap.add_argument("--synthetic", type=int, default=0, help="use N synthetic MNIST-shaped samples instead of the dataset files")
    ap.add_argument("--checkpoint-dir", default=None, help="chief-only TF-format checkpoints + TensorBoard events")
    return ap.parse_args()


def make_strategy(a: argparse.Namespace):
    devices = [a.device] if a.device else None
    if a.mirrored:
        return tfdl.distribute.MirroredStrategy(devices=devices)
    tf_config = a.tf_config
    if a.workers:
        tf_config = json.dumps({"cluster": {"worker": a.workers.split(",")}, "task": {"type": "worker", "index": a.index}})
    comm = tfdl.distribute.experimental.CollectiveCommunication[a.communication]
    return tfdl.distribute.experimental.MultiWorkerMirroredStrategy(comm, tf_config=tf_config, devices=devices)


def unbatched_training_set(synthetic: int):
    kw = {"synthetic_size": (synthetic, 64)} if synthetic else {}
    splits, _info = tfdl.datasets.load(name="mnist", with_info=True, as_supervised=True, **kw)

    def to_unit_range(image, label):
        return tfdl.utils.cast(image, tfdl.utils.float32) / 255, label

    return splits["train"].map(to_unit_range).cache().shuffle(SHUFFLE_BUFFER)


def run_evaluator(a: argparse.Namespace) -> None:
    """The `evaluator` task of the cluster (README.md:57 of the reference): outside the training group, it evaluates every
    checkpoint the chief writes to --checkpoint-dir on the test split."""
    if not a.checkpoint_dir:
        raise SystemExit("an evaluator task needs --checkpoint-dir (where the chief saves)")
    kw = {"synthetic_size": (a.synthetic, max(64, a.synthetic // 8))} if a.synthetic else {}
    splits, _ = tfdl.datasets.load(name="mnist", with_info=True, as_supervised=True, **kw)
    test = splits["test"].map(lambda im, lb: (tfdl.utils.cast(im, tfdl.utils.float32) / 255, lb)).batch(PER_WORKER_BATCH)
    model = tfdl.models.build_and_compile_cnn_model(learning_rate=a.learning_rate)
    ev = tfdl.keras.utils.SidecarEvaluator(model, test, a.checkpoint_dir, until_step=a.epochs * a.steps_per_epoch,
                                           log_dir=os.path.join(a.checkpoint_dir, "logs"), poll_interval_s=0.2, timeout_s=300)
    for out in ev.start():
        print(json.dumps({"evaluator": True, **{k: v for k, v in out.items()}}))


def main() -> None:
    a = parse_args()
    tfdl.datasets.disable_progress_bar()
    if not a.mirrored and tfdl.cluster.TFConfigClusterResolver.from_env(a.tf_config).task_type == "evaluator":
        return run_evaluator(a)
    strategy = make_strategy(a)  # TF_CONFIG is read here, before anything else touches the devices
    global_batch = PER_WORKER_BATCH * strategy.num_replicas_in_sync

    options = tfdl.data.Options()
    options.experimental_distribute.auto_shard_policy = tfdl.data.experimental.AutoShardPolicy.OFF
    train = unbatched_training_set(a.synthetic).batch(global_batch).with_options(options)

    with strategy.scope():  # variables created here are mirrored on every replica
        model = tfdl.models.build_and_compile_cnn_model(learning_rate=a.learning_rate)

    callbacks = []
    if a.checkpoint_dir:  # only the chief writes (README.md:51 of the reference)
        callbacks = [tfdl.keras.callbacks.ModelCheckpoint(os.path.join(a.checkpoint_dir, "ckpt-{epoch}"), save_weights_only=True),
                     tfdl.keras.callbacks.TensorBoard(os.path.join(a.checkpoint_dir, "logs")),
                     tfdl.keras.callbacks.BackupAndRestore(os.path.join(a.checkpoint_dir, "backup"))]
    history = model.fit(x=train, epochs=a.epochs, steps_per_epoch=a.steps_per_epoch, callbacks=callbacks,
                        verbose=1 if strategy.is_chief else 0)
    if strategy.is_chief:
        print(json.dumps({"replicas": strategy.num_replicas_in_sync, "engine": model.engine_name,
                          "final_loss": history.history["loss"][-1],
                          "final_accuracy": history.history.get("sparse_categorical_accuracy", [None])[-1]}))
    strategy.close()


if __name__ == "__main__":
    main()
