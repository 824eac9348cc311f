"""Command-line tools (reference: S/models/utils/{DistriOptimizerPerf,LocalOptimizerPerf,ImageNetSeqFileGenerator,
COCOSeqFileGenerator}.scala, S/nn/mkldnn/Perf.scala, S/utils/ConvertModel.scala). Run as
``python -m bigdl_b200.cli.<tool> --help``; multi-GPU tools are launched with torchrun (scripts/bigdl-b200-run.sh)."""
