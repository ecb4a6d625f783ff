"""B200-native BSRNN separation path (drop-in for phhusson/SpeechSeparation's bsrnn.BSRNN).

Only what the hot path needs lives here: ``csrc/`` (CUDA kernels + C ABI), the ``BSRNN``
module mirror, the signal wrap and the row-sharding helper for multi-GPU runs.
"""
from .bsrnn import (BSRNN, BandwiseLSTM, NormRNNResidual, TimewiseLSTM, TrainableConstantModule,
                    band_features, generate_bandsplits, merge_channels)
from .wrap import (BatchSeparator, StreamingSeparator, infer, metrics_from_stats, pair_stats, stft_pack, train_infer,
                   unpack_istft)
from .shard import shard_range, aggregate_throughput
from .export import ToTrace, trace_streaming_step          # registers the bsrnn_b200:: operators

__all__ = ["BSRNN", "BandwiseLSTM", "TimewiseLSTM", "NormRNNResidual", "TrainableConstantModule",
           "generate_bandsplits", "band_features", "merge_channels", "stft_pack", "unpack_istft", "infer",
           "StreamingSeparator", "BatchSeparator", "pair_stats", "metrics_from_stats", "train_infer", "shard_range", "aggregate_throughput",
           "ToTrace", "trace_streaming_step"]
