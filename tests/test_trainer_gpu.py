"""The trainers and the eval loop end to end on the GPU, through their `main()`:

  * resnet_cifar_train: a run that is KILLED (SIGKILL) after it has written a checkpoint and is then
    restarted on the same --train_dir continues bit-identically to a run that was never interrupted
    (reference: MonitoredTrainingSession restores from checkpoint_dir, resnet_cifar_train.py:330-344);
    the chief writes TensorBoard events with the reference's four scalars (:275-287).
  * --mode eval (resnet_cifar_eval.evaluate, reference resnet_cifar_eval.py:60-143) on the reference's
    REAL trained weights, loaded from a frozen GraphDef .pb through tf_io: precision equals the CPU
    oracle's precision on the same records; Precision / Best Precision events are written.
  * checkpoints in TF-V2 bundle format round-trip through the trainer.
"""
import os
import signal
import subprocess
import sys
import time

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

from distributed_tensorflow_resnet_b200 import FLAGS, resnet_cifar_train as ct, tf_io  # noqa: E402
from distributed_tensorflow_resnet_b200.train_utils import read_event_scalars, latest_checkpoint  # noqa: E402

COMMON = ["--train_data_path", "synthetic:2048", "--batch_size", "32", "--resnet_size", "8", "--compute_dtype", "bf16",
          "--save_checkpoint_steps", "10", "--seed", "7"]


def _run(train_dir, steps, extra=()):
    FLAGS.reset()
    m = ct.main(COMMON + ["--train_dir", str(train_dir), "--train_steps", str(steps)] + list(extra))
    torch.cuda.synchronize()
    return m


def test_killed_trainer_resumes_bit_identically(tmp_path):
    a, b, logs = tmp_path / "a", tmp_path / "b", tmp_path / "logs"
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    cmd = [sys.executable, "-m", "distributed_tensorflow_resnet_b200.resnet_cifar_train"] + COMMON + \
          ["--train_dir", str(a), "--train_steps", "1000000", "--log_dir", str(logs)]
    proc = subprocess.Popen(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, cwd=ROOT)
    try:
        t0 = time.time()
        while not (a / "model.ckpt-20.npz").exists():
            assert proc.poll() is None, proc.stdout.read().decode()[-3000:]
            assert time.time() - t0 < 300, "trainer did not reach step 20 in 5 minutes"
            time.sleep(0.05)
    finally:
        os.kill(proc.pid, signal.SIGKILL)         # the exact PID we started; no clean shutdown, no final checkpoint
        proc.wait()
    path = latest_checkpoint(str(a))
    s0 = int(os.path.basename(path).split("-")[1].split(".")[0])
    assert s0 >= 20 and s0 % 10 == 0
    z = np.load(path)
    assert int(z["global_step"]) == s0
    # the chief wrote TensorBoard events with the reference's scalars before it was killed
    tags = {t for _, t, _ in read_event_scalars(str(logs))}
    assert {"cross_entropy", "cost", "learning_rate", "Precision"} <= tags
    target = s0 + 15
    m1 = _run(a, target)                            # resumes from step s0
    assert m1.global_step == target
    w1, mom1, s1 = m1.executor.W.clone(), m1.executor.Mom.clone(), m1.executor.S.clone()
    m2 = _run(b, target)                            # never interrupted
    assert m2.global_step == target
    assert torch.equal(w1, m2.executor.W) and torch.equal(mom1, m2.executor.Mom) and torch.equal(s1, m2.executor.S)
    # a third start on the finished directory has nothing left to do and changes nothing
    m3 = _run(a, target)
    assert m3.global_step == target and torch.equal(m3.executor.W, w1)


def test_tf_bundle_checkpoints_through_the_trainer(tmp_path):
    m = _run(tmp_path, 12, ["--checkpoint_format", "tf"])
    prefix = latest_checkpoint(str(tmp_path))
    assert prefix.endswith("model.ckpt-12") and os.path.exists(prefix + ".index")
    idx = tf_io.read_checkpoint_index(prefix)
    names = {v.name for v in m.graph.variables.values()}
    assert names <= set(idx) and "conv2d/kernel/Momentum" in idx and idx["global_step"]["shape"] == []
    sd = tf_io.read_checkpoint(prefix)
    assert np.array_equal(sd["conv2d/kernel"], m.executor.var_view("conv2d/kernel").cpu().numpy())
    w = m.executor.W.clone()
    m2 = _run(tmp_path, 12, ["--checkpoint_format", "tf"])          # restore only
    assert m2.global_step == 12 and torch.equal(m2.executor.W, w)


def test_eval_loop_on_the_reference_trained_weights(tmp_path):
    from oracle import fixtures, resnet_oracle as ro, cifar_oracle as co
    w = fixtures.frozen_weights()
    train_dir, eval_dir = tmp_path / "train", tmp_path / "eval"
    os.makedirs(train_dir)
    tf_io.write_frozen_graph(str(train_dir / "resnet50_cifar_frozen_model_eval.pb"), w)
    (train_dir / "checkpoint").write_text('model_checkpoint_path: "resnet50_cifar_frozen_model_eval.pb"\n')
    FLAGS.reset()
    res = ct.main(["--mode", "eval", "--train_dir", str(train_dir), "--eval_dir", str(eval_dir), "--eval_data_path",
                   "synthetic:300", "--eval_batch_count", "2", "--eval_once", "true", "--compute_dtype", "f32",
                   "--resnet_size", "50", "--seed", "3"])
    assert res["total"] == 200
    rec = co.synthetic_records(300, "cifar10", 3)
    img, onehot, lab, _ = co.decode(rec, np.arange(200), train=False, pad=2)
    net = ro.OfficialResNetV2("cifar10", 50, 10, dtype=torch.float64, params=w)
    lg = net.forward(torch.from_numpy(img).double(), False)
    want = float((lg.val.argmax(1).numpy() == lab).mean())
    assert abs(res["precision"] - want) < 1e-9 and res["best_precision"] == res["precision"]
    tags = {t for _, t, _ in read_event_scalars(str(eval_dir))}
    assert {"Precision", "Best Precision"} <= tags


def test_imagenet_trainer_runs_on_the_preprocessing_pipeline(tmp_path):
    """resnet_imagenet_train.main on the synthetic ImageNet input (uint8 images of random sizes -> rn_vgg_preprocess):
    warm-up learning rates of the reference's hook (:236-260), loss finite, checkpoint + resume."""
    import io
    from distributed_tensorflow_resnet_b200 import resnet_imagenet_train as it
    FLAGS.reset()
    args = ["--train_data_path", "synthetic:16", "--batch_size", "4", "--resnet_size", "18", "--train_steps", "6",
            "--train_dir", str(tmp_path), "--compute_dtype", "bf16", "--seed", "2"]
    m = it.main(args)
    torch.cuda.synchronize()
    assert m.global_step == 6 and np.isfinite(m.cost)
    assert abs(m.lrn_rate - it.imagenet_learning_rate(5)) < 1e-12           # rate for step 6 was set from step 5
    pipe = m._images._rn_pipeline
    assert pipe.draws.cpu().numpy()[:, 0].min() >= 256
    w = m.executor.W.clone()
    FLAGS.reset()
    m2 = it.main(args)                                                     # finished directory: restore only
    assert m2.global_step == 6 and torch.equal(m2.executor.W, w)
    # the ImageNet eval loop (resnet_imagenet_eval.py:169-230) on that checkpoint: vgg eval preprocessing (central crop),
    # BN on the moving statistics, Precision / Best Precision; through the combined entry point resnet_imagenet_main
    from distributed_tensorflow_resnet_b200 import resnet_imagenet_main, resnet_imagenet_horovod_train
    from distributed_tensorflow_resnet_b200.train_utils import read_event_scalars
    FLAGS.reset()
    eval_dir = tmp_path / "eval"
    res = resnet_imagenet_main.main(["--mode", "eval", "--eval_data_path", "synthetic:16", "--batch_size", "4",
                                     "--resnet_size", "18", "--train_dir", str(tmp_path), "--eval_dir", str(eval_dir),
                                     "--eval_batch_count", "3", "--eval_once", "true", "--seed", "2"])
    assert res["global_step"] == 6 and res["total"] == 12 and 0.0 <= res["precision"] <= 1.0 and np.isfinite(res["loss"])
    assert {"Precision", "Best Precision"} <= {t for _, t, _ in read_event_scalars(str(eval_dir))}
    # the Horovod-named trainer is the same loop with the all-reduce flags preset; a finished directory only restores
    FLAGS.reset()
    m3 = resnet_imagenet_horovod_train.main(args)
    assert FLAGS.use_horovod and FLAGS.variable_update == "horovod" and m3.global_step == 6


def test_resnet_single_entry_point_trains_the_north_star_builders(tmp_path):
    """resnet_single.main: the upstream builders behind the nine-field HParams of resnet_single.py:194-202 (5 residual
    units per stage = ResNet-32, leaky ReLU 0.1, weight decay on the DW variables), checkpoints under log_root/train."""
    from distributed_tensorflow_resnet_b200 import resnet_single
    FLAGS.reset()
    m = resnet_single.main(["--train_data_path", "synthetic:512", "--train_steps", "4", "--log_root", str(tmp_path),
                            "--batch_size", "16", "--seed", "1"])
    torch.cuda.synchronize()
    assert m.global_step == 4 and np.isfinite(m.cost)
    assert m.hps.num_residual_units == 5 and m.hps.relu_leakiness == 0.1
    names = list(m.graph.variables)
    assert any(n.endswith("/DW") for n in names) and any("unit_3_4" in n for n in names)      # upstream naming, 5 units per stage
    assert (tmp_path / "train").is_dir() and any(f.startswith("model.ckpt-4") for f in os.listdir(tmp_path / "train"))


def test_freeze_then_predict_from_pb(tmp_path):
    """resnet_cifar_frozen_model.main freezes the newest checkpoint into a GraphDef of Const nodes (and predicts one
    batch); resnet_cifar_predict_from_pd.main loads that .pb into a fresh eval graph and must print the very same
    predictions — the round trip the reference's two scripts make through tf.import_graph_def."""
    import io
    from distributed_tensorflow_resnet_b200 import resnet_cifar_frozen_model as fz, resnet_cifar_predict_from_pd as pp
    FLAGS.reset()
    ct.main(["--train_data_path", "synthetic:1024", "--batch_size", "32", "--resnet_size", "8", "--train_steps", "5",
             "--train_dir", str(tmp_path), "--seed", "4"])
    FLAGS.reset()
    buf = io.StringIO()
    a = fz.main(["--train_dir", str(tmp_path), "--resnet_size", "8", "--batch_size", "50", "--seed", "4"], out=buf)
    assert os.path.exists(a["path"]) and "The precision:" in buf.getvalue()
    got = tf_io.read_frozen_graph(a["path"])
    assert len(got) == a["tensors"] and any(k.endswith("moving_variance") for k in got)
    FLAGS.reset()
    b = pp.main(["--frozen_model", a["path"], "--resnet_size", "8", "--EVAL_NUM", "50", "--seed", "4"], out=io.StringIO())
    assert b["loaded"] == a["tensors"]
    assert np.array_equal(a["prediction"], b["prediction"]) and np.array_equal(a["label"], b["label"])
    assert a["precision"] == b["precision"]
