"""chaq-sdfgen_b200 -- B200-native signed-distance-field generation behind the interface of
chaquator/chaq-sdfgen's OpenMP program (reference openmp/df.h:6 and openmp/sdfgen.c:261-300).

The product is csrc/ (hand-written sm_100a CUDA kernels + the C ABI of include/chaqsdf.h, built in-tree as
libchaqsdf_b200.so). This Python layer is a thin ctypes mirror of that ABI for tests and bench.py; torch is
used only for device buffers, streams and torch.distributed. There is no CPU fallback: loading fails loudly
if the library is missing, and every compute call fails without a CUDA device.

The directory name has a hyphen, so import it through `load_package()` in __graft_entry__.py (importlib).
"""
from .sdfgen import (ChaqSdfError, LIB_PATH, build_library, dist_transform_2d, lib, sdf_u8, sdf_u8_multi, sdf_u8_batched_device,  # noqa: F401
                     sdf_u8_device, sdf_u8_device_static, STATIC_SCRATCH_BYTES, sdf_fields_device, rowscan_device, envelope_device, workspace_bytes,
                     launch_count, profile_enable, profile_mean_ms, reload_tuning, sdf_u8_sharded, inter_row_bytes, shard_rows, halo_plan, exchange_halos, halo_pull_list, pull_halos, pipeline_plan, col_part, rowscan_scatter_device, fields_from_inter_device, sdf_fields_a2a, transpose_u8_device, envelope_cols_device, sdf_u8_tband_device, tband_setup, tband_input_view, sdf_u8_sharded_t, atlas_pack, glyph_boxes_device, atlas_blit_device, atlas_u8, SYMBOLS)
