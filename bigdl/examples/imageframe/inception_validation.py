"""pyspark/bigdl/examples/imageframe/inception_validation.py: top-1 accuracy of a saved model on ImageNet validation
SequenceFiles through the ImageFrame pipeline.   python -m bigdl.examples.imageframe.inception_validation <seqFolder> <model> <batch>"""
import os
import sys


def get_data(url, sc=None, data_type="val"):
    from bigdl_b200.transform.vision.image import SeqFileFolder
    return SeqFileFolder.files_to_image_frame(url=os.path.join(url, data_type), sc=sc, class_num=1000)


def run(image_path, model_path, batch_size):
    from bigdl_b200.nn.module import Module
    from bigdl_b200.optim.validation import Top1Accuracy
    from bigdl_b200.transform.vision.image import (CenterCrop, ChannelNormalize, ImageFrameToSample, MatToTensor, Pipeline,
                                                   PixelBytesToMat, Resize)
    from bigdl_b200.utils.engine import Engine

    Engine.init()
    transformer = Pipeline([PixelBytesToMat(), Resize(256, 256), CenterCrop(224, 224),
                            ChannelNormalize(123.0, 117.0, 104.0), MatToTensor(),
                            ImageFrameToSample(input_keys=["imageTensor"], target_keys=["label"])])
    frame = transformer(get_data(image_path))
    model = Module.loadModule(model_path)
    result = model.evaluate(frame, int(batch_size), [Top1Accuracy()])
    print("top1 accuracy", result[0])
    return result


if __name__ == "__main__":
    if len(sys.argv) != 4:
        print("parameters needed : <imagePath> <modelPath> <batchSize>")
        sys.exit(1)
    run(sys.argv[1], sys.argv[2], sys.argv[3])
