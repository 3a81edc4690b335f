"""neural_sound_generation_b200 -- B200-native (sm_100a) VQ-VAE hot path.

Drop-in for the VQ-VAE path of dendisuhubdy/neural_sound_generation: the modules below keep the
reference's names and call signatures (``vector_quantization.vq / vq_st``, ``models.VQEmbedding /
ResBlock / VQVAE / weights_init / to_scalar``, ``train.train_vqvae``, ``test.test_vqvae``) and run on
hand-written CUDA kernels behind the C ABI in ``include/nsg_b200.h``.  There is no CPU fallback.
"""
from . import _cabi  # noqa: F401  (does not load the library until first use)
from .vector_quantization import vq, vq_st  # noqa: F401
from .models import VQEmbedding, ResBlock, VQVAE, weights_init, to_scalar  # noqa: F401
from .train import train_vqvae  # noqa: F401
from .test import test_vqvae  # noqa: F401
from . import optim  # noqa: F401
from .graph import GraphedTrainStep, GraphedInference  # noqa: F401
from . import dist  # noqa: F401  (DataParallelStep: batch-sharded data parallelism)

VectorQuantizer = VQEmbedding   # alias for the name BASELINE.json's north_star uses
__version__ = '0.1'
from . import data  # noqa: F401  (device-side mel batch assembly: MelNpyDataset, DeviceMelLoader)
from . import audio  # noqa: F401  (mel <-> audio: STFT / ISTFT / Griffin-Lim kernels, audio_tacotron.py names)
from .models_n4 import VAE, GatedActivation, GatedMaskedConv2d, GatedPixelCNN  # noqa: F401  (row N4: the other model families)
from . import tiled  # noqa: F401  (W-tiled inference of one long utterance, one tile per GPU)
