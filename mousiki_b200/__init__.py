"""mousiki_b200 — B200-native batched Opus decoder (drop-in for mousiki's packet-decode path).

The package is a thin host binding over libmousiki_b200.so (hand-written CUDA for sm_100a, C ABI in
include/mousiki_b200.h). See DESIGN.md for the kernels and INTEGRATION.md for the Rust-side binding.
"""
from .batch_decoder import (BatchDecoder, DecoderError, EXPORTS, LiteBatchDecoder, PacketError, QUIRKS_MOUSIKI, load_library, pack_packets,  # noqa: F401
                            packet_get_bandwidth, packet_get_mode, packet_get_nb_channels, packet_get_nb_frames, packet_get_nb_samples,
                            packet_get_samples_per_frame)

__all__ = ["BatchDecoder", "DecoderError", "EXPORTS", "LiteBatchDecoder", "PacketError", "QUIRKS_MOUSIKI", "load_library", "pack_packets", "packet_get_bandwidth",
           "packet_get_mode", "packet_get_nb_channels", "packet_get_nb_frames", "packet_get_nb_samples", "packet_get_samples_per_frame"]
