"""``MagicDriveSTDiT3Config`` — constructor surface of the drop-in.

Mirrors the keyword set of the reference config (magicdrive_stdit3.py:361-472).  Flags that select
third-party kernels in the reference (``enable_flash_attn``, ``enable_xformers``,
``enable_layernorm_kernel``) are accepted and ignored: there is exactly one execution path here,
the sm_100a one.  Unknown keys (``frame_emb_cls`` / ``frame_emb_param``) ride on
``PretrainedConfig(**kwargs)`` exactly as they do upstream.
"""
from transformers import PretrainedConfig

_DEFAULTS = dict(
    input_size=(1, 32, 32), input_sq_size=512, force_pad_h_for_sp_size=None, simulate_sp_size=[],
    in_channels=4, patch_size=(1, 2, 2), hidden_size=1152, depth=28, num_heads=16, mlp_ratio=4.0,
    class_dropout_prob=0.1, pred_sigma=True, drop_path=0.0, caption_channels=4096,
    model_max_length=300, qk_norm=True, enable_flash_attn=False, enable_xformers=False,
    enable_layernorm_kernel=False, enable_sequence_parallelism=False, freeze_y_embedder=False,
    with_temp_block=True, freeze_x_embedder=False, freeze_old_embedder=False,
    freeze_temporal_blocks=False, freeze_old_params=False, zero_and_train_embedder=None,
    only_train_base_blocks=False, only_train_temp_blocks=False, only_train_extra_blocks=False,
    qk_norm_trainable=False, sequence_parallelism_temporal=False, control_depth=13,
    use_x_control_embedder=False, use_st_cross_attn=False, uncond_cam_in_dim=(3, 7),
    cam_encoder_cls=None, cam_encoder_param={}, bbox_embedder_cls=None, bbox_embedder_param={},
    map_embedder_cls=None, map_embedder_param={}, map_embedder_downsample_rate=4,
    micro_frame_size=17, control_skip_cross_view=True, control_skip_temporal=True,
    brushnet_skip_cross_attn=True, use_lora_base_blocks=False, lora_rank=16, lora_alpha=32,
    lora_dropout=0.0,
)


class MagicDriveSTDiT3Config(PretrainedConfig):
    model_type = "MagicDriveSTDiT3"

    def __init__(self, **kwargs):
        for key, default in _DEFAULTS.items():
            value = kwargs.pop(key, default)
            if isinstance(default, (dict, list)) and value is default:
                value = type(default)(default)  # no shared mutable defaults
            setattr(self, key, value)
        super().__init__(**kwargs)
