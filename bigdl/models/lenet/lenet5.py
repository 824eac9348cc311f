"""P/models/lenet/lenet5.py: ``build_model(class_num)`` and a train/test CLI."""
import argparse

from bigdl_b200.models.lenet import LeNet5


def build_model(class_num):
    return LeNet5(class_num)


def main(argv=None):
    from bigdl.models.lenet.utils import preprocess_mnist
    from bigdl.nn.criterion import ClassNLLCriterion
    from bigdl.optim.optimizer import EveryEpoch, MaxEpoch, Optimizer, SGD, Top1Accuracy

    ap = argparse.ArgumentParser()
    ap.add_argument("-a", "--action", default="train")
    ap.add_argument("-b", "--batchSize", type=int, default=128)
    ap.add_argument("-n", "--endTriggerNum", type=int, default=20)
    ap.add_argument("-d", "--dataPath", default="/tmp/mnist")
    ap.add_argument("--checkpointPath", default=None)
    a = ap.parse_args(argv)
    train, test = preprocess_mnist(None, a, a.dataPath)
    model = build_model(10)
    if a.action == "train":
        opt = Optimizer(model=model, training_set=train, criterion=ClassNLLCriterion(),
                        optim_method=SGD(learningrate=0.01, learningrate_decay=0.0002),
                        end_trigger=MaxEpoch(a.endTriggerNum), batch_size=a.batchSize)
        opt.setValidation(EveryEpoch(), test, [Top1Accuracy()], a.batchSize)
        if a.checkpointPath:
            opt.setCheckpoint(a.checkpointPath, EveryEpoch())
        opt.optimize()
    res = model.evaluate(test, a.batchSize, [Top1Accuracy()])
    print(res)


if __name__ == "__main__":
    main()
