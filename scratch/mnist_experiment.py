"""The reference's own experiment, end to end, on the GPU backend: MNISTTrainer.main
(src/MNISTTrainer.java:221-245) = load the four IDX files, train, testOnTrainingData, testOnTestData, logTest -- for the
three schedules the reference logged (logs/trainLog.csv:3-5) plus BASELINE configs[2] with C1's hyper-parameters.

  python scratch/mnist_experiment.py [--out gpurun_out/mnist] [--skip-500k]

The image files are the deterministic MNIST-shaped SURROGATE (graph-neural-net_b200/mnist.py:write_surrogate_images,
generated from the reference's REAL label files, tests/golden/mnist): the real image files are absent from the reference
mount, so every number printed here is "surrogate, not MNIST".  Per epoch (every 60000 / B steps) the mean loss over the
first 601 rows is recorded (validate(601), T:65, T:102-113: the reference defines no per-epoch loss; SURVEY 8d does).
Beside each wall time stands the CPU port's extrapolated time: cpu_samples_per_s measured by bench.py's cpu_baseline
(the oracle's faithful loops, 1 thread) x iterations x B.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge  # noqa: E402

SCHEDULES = [
    # name, class, dims, iterations, step, B, reference row
    ("c1_100k", "softmax", [784, 100, 50, 10], 100000, 0.01, 8, "logs/trainLog.csv:3 (0.98598 / 0.96680 on MNIST)"),
    ("c2_100k", "softmax", [784, 300, 100, 10], 100000, 0.0125, 16, "logs/trainLog.csv:4 = main's literals M:224-230 (0.99973 / 0.97800 on MNIST)"),
    ("c3_100k", "general", [784, 100, 50, 10], 100000, 0.01, 8, "BASELINE configs[2]: sigmoid / half squared error, C1's hyper-parameters (no reference row)"),
    ("c2_500k", "softmax", [784, 300, 100, 10], 500000, 0.0125, 16, "logs/trainLog.csv:5 (1.00000 / 0.97960 on MNIST)"),
]
CPU_SAMPLES_PER_S = {"c1": 3297.0, "c2": 1036.0, "c3": 1798.0}   # bench.py cpu_baseline on the GPU box (profiles/r02_*)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "mnist"))
    ap.add_argument("--skip-500k", action="store_true")
    args = ap.parse_args()
    g = ge.load_package()
    from oracle import oracle as orc   # the checker formats the CSV row the way M:211-219 does
    os.makedirs(args.out, exist_ok=True)
    data_dir = os.path.join(args.out, "data")
    t0 = time.time()
    g.mnist.ensure_surrogate(data_dir, os.path.join(ROOT, "tests", "golden", "mnist"))
    gen_s = time.time() - t0
    results = []
    for name, cls, dims, iters, step, B, refrow in SCHEDULES:
        if args.skip_500k and iters > 100000:
            continue
        log_dir = os.path.join(args.out, "logs_" + name)
        w0 = time.time()
        if cls == "softmax":
            net = g.SoftmaxCrossEntropyNeuralNet(dims, g.LEAKY_RELU, g.LEAKY_RELU_PRIME)      # M:232-235
        else:
            net = g.GeneralNeuralNet(dims, g.SIGMOID, g.SIGMOID_PRIME, g.SIGMOID, g.SIGMOID_PRIME,
                                     g.HALF_SQUARED_ERROR, g.HALF_SQUARED_ERROR_PRIME)
        tr = g.MNISTTrainer(net, data_dir, log_dir)                                            # M:236 (loads the IDX files)
        load_s = time.time() - w0
        epoch = 60000 // B
        per_epoch = []
        t1 = time.time()
        done = 0
        while done < iters:                                                                    # M:239 in epoch-sized chunks
            n = min(epoch, iters - done)
            tr.trainer.train(n, step, B, 0.9, False, None, None)
            done += n
            per_epoch.append((done, tr.trainer.validate(601)))
        net.sync()
        train_s = time.time() - t1
        t2 = time.time()
        train_acc = tr.testOnTrainingData()                                                    # M:241, train first
        test_acc = tr.testOnTestData()
        eval_s = time.time() - t2
        tr.logTest(dims, iters, step, B, 0.9, False, train_acc, test_acc)
        wall = time.time() - w0
        row = open(os.path.join(log_dir, "trainLog.csv")).read().strip().splitlines()[-1]
        cpu_s = iters * B / CPU_SAMPLES_PER_S[name[:2]]
        r = {"name": name, "dims": dims, "iterations": iters, "step": step, "batch": B, "reference_row": refrow,
             "csv_row": row, "train_acc": train_acc, "test_acc": test_acc, "load_s": load_s, "train_s": train_s,
             "eval_s": eval_s, "main_wall_s": wall, "us_per_step": train_s / iters * 1e6,
             "cpu_port_extrapolated_train_s": cpu_s, "per_epoch_validate601": per_epoch, "data": "surrogate, not MNIST"}
        results.append(r)
        print("%s: %s | load %.1f s, train %.2f s (%.1f us/step incl. per-epoch validate), eval %.2f s, whole main %.1f s; "
              "CPU port (1 thread) would train for %.0f s" % (name, row, load_s, train_s, r["us_per_step"], eval_s, wall,
                                                               cpu_s), flush=True)
        print("   per-epoch validate(601): " + " ".join("%.4g" % l for _, l in per_epoch[:12]) +
              (" ... %.4g" % per_epoch[-1][1] if len(per_epoch) > 12 else ""), flush=True)
        assert row == orc.format_log_row(dims, iters, step, B, 0.9, False, train_acc, test_acc).strip()
        tr.trainingLog.close()
        tr.trainer.dataMaster.close()
        tr.testingPartition.close()
        net.close()
    with open(os.path.join(args.out, "mnist_surrogate.json"), "w") as f:
        json.dump({"surrogate_generation_s": gen_s, "runs": results}, f, indent=1)


if __name__ == "__main__":
    main()
