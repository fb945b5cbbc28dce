"""vit-research_b200 — B200-native ViT embed + k-NN vote hot path of Gurinder-Vasanta/vit-research.

Import as `vit_research_b200` (the hyphenated directory is exposed by the shim package of that
name at the repo root).  See DESIGN.md for the path, its boundary and the kernels.
"""
from ._lib import VitkError, build, load  # noqa: F401
from .embed import (ViTEmbedder, ViTGeometry, extract_embeddings, hf_vit_embed_batch,  # noqa: F401
                    install, random_init_state_dict)
from . import knn  # noqa: F401
from .knn import (BankClient, BankCollection, BankIndex, FrameRetriever, RattChunkRetriever,  # noqa: F401
                  determine_class_pass1, determine_class_pass2, load_label_banks, temp_smax,
                  LABEL_NAMES, determine_class, classify_stream_with_upserts, match_where,
                  one_hot_metadata, vote_from_counts)
from . import hmm as hmm_module  # noqa: F401,E402
from .hmm import ViterbiStream, hmm, viterbi_from_counts, viterbi_from_probs  # noqa: F401,E402
from . import store  # noqa: F401,E402
from .chunk_encoder import ChunkEncoder  # noqa: F401,E402
from . import shard  # noqa: F401,E402
from .shard import (frame_ranges, gather_frames, label_and_smooth_stream, label_frames_sharded,
                    round_robin_chunks)  # noqa: F401,E402
