"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/imx.h declares, argument validation works without a GPU, the host-side modules keep
the reference's names / keys, and CPU tensors are rejected (no fallback)."""
import ctypes
import os
import re

import pytest
import torch

from tests import helpers as H

ROOT = H.ROOT


def test_library_exports_every_declared_symbol():
    from imaginaire_b200 import _C
    decls = _C.parse_header()
    header = open(os.path.join(ROOT, 'include', 'imx.h')).read()
    declared = set(re.findall(r'IMX_API[^;(]*?\b(imx_\w+)\s*\(', header))
    assert declared == set(decls) and len(declared) >= 20
    dll = ctypes.CDLL(_C.LIB_PATH)
    for name in declared:
        assert hasattr(dll, name), name
    assert _C.lib().raw('imx_version')() >= 100


def test_argument_validation_needs_no_gpu():
    from imaginaire_b200 import _C
    lib = _C.lib()
    # kernel_size != 1 is rejected before anything touches the device
    rc = lib.raw('imx_resample2d_fwd_f32')(1, 1, 1, 1, 3, 8, 8, 8, 8, 3, None)
    assert rc != 0 and 'kernel_size' in lib.last_error()
    rc = lib.raw('imx_bn_stats_nhwc_bf16')(1, 1, 1, 1, 16, 12, 0, None)
    assert rc != 0 and 'multiple of 8' in lib.last_error()
    assert lib.raw('imx_conv2d_wgrad_workspace_bytes')(8, 128, 256, 128, 128, 512, 512, 5, 5, 1, 2) > 0
    with pytest.raises(_C.ImxError):
        lib.call('imx_conv2d_igemm_fprop_bf16', 1, 1, None, None, 1, None, 1, 8, 8, 12, 8, 8, 3, 3, 1, 1, 0, None)


def test_cpu_tensors_are_rejected():
    from imaginaire_b200 import ops
    from imaginaire_b200.model_utils.fs_vid2vid import resample
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        ops.resample2d_fwd(torch.zeros(1, 1, 4, 4), torch.zeros(1, 2, 4, 4))
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        resample(torch.zeros(1, 1, 4, 4), torch.zeros(1, 2, 4, 4))
    from imaginaire_b200.layers import Conv2dBlock
    with pytest.raises(RuntimeError, match='no CPU fallback'):
        Conv2dBlock(8, 8, 3, padding=1)(torch.zeros(1, 8, 4, 4))


def test_generator_state_dict_keys_match_reference():
    d = H.load_golden('generator_unit_test.npz')
    _, net = H.build_generator('spade_unit_test.yaml')
    sd = net.state_dict()
    assert list(sd.keys()) == [str(k) for k in d['keys']]
    assert len(sd) == 263
    assert H.state_dict_digest(sd) == str(d['state_dict_sha256'])


def test_block_api_surface():
    import types
    from imaginaire_b200.layers import Conv2dBlock, Res2dBlock, SpatiallyAdaptiveNorm
    an = types.SimpleNamespace(num_filters=8, kernel_size=3, separate_projection=True, cond_dims=5,
                               weight_norm_type='spectral', activation_norm_type='sync_batch',
                               activation_norm_params=types.SimpleNamespace(affine=False))
    blk = Res2dBlock(16, 8, activation_norm_type='spatially_adaptive', activation_norm_params=an,
                     weight_norm_type='spectral', bias=[True, True, False], order='NACNAC')
    assert blk.learn_shortcut and blk.conditional
    assert list(blk.conv_block_0.layers.keys()) == ['norm', 'nonlinearity', 'conv']
    assert list(blk.conv_block_s.layers.keys()) == ['norm', 'conv']
    keys = set(blk.state_dict().keys())
    for k in ('conv_block_0.layers.conv.weight_orig', 'conv_block_0.layers.conv.weight_u',
              'conv_block_0.layers.conv.weight_v', 'conv_block_0.layers.conv.bias',
              'conv_block_0.layers.norm.norm.running_mean', 'conv_block_0.layers.norm.norm.weight',
              'conv_block_0.layers.norm.mlps.0.0.layers.conv.weight_orig',
              'conv_block_0.layers.norm.gammas.0.layers.conv.bias',
              'conv_block_0.layers.norm.betas.0.layers.conv.weight_v'):
        assert k in keys, k
    assert 'conv_block_s.layers.conv.bias' not in keys
    norm = blk.conv_block_0.layers['norm']
    assert isinstance(norm, SpatiallyAdaptiveNorm) and norm.conditional
    # sync_batch always owns weight / bias; `affine` only decides whether they train
    assert not norm.norm.weight.requires_grad
    blk2 = Conv2dBlock(4, 4, 3, padding=1, nonlinearity='leakyrelu', order='CNA')
    assert list(blk2.layers.keys()) == ['conv', 'nonlinearity'] and not blk2.conditional


def test_config_loader():
    from imaginaire_b200.config import Config
    from imaginaire_b200.utils.data import (get_crop_h_w, get_paired_input_image_channel_number,
                                            get_paired_input_label_channel_number)
    cfg = Config(os.path.join(ROOT, 'configs', 'spade_c3_256x512.yaml'))
    assert get_paired_input_label_channel_number(cfg.data) == 35
    assert get_paired_input_image_channel_number(cfg.data) == 3
    assert get_crop_h_w(cfg.data.train.augmentations) == (256, 512)
    assert cfg.gen.activation_norm_params.separate_projection is True
    assert cfg.gen_opt.lr == 1e-4 and cfg.trainer.model_average is True


@pytest.mark.skipif(not os.path.isdir('/root/reference/configs'), reason='reference tree not mounted')
def test_reference_yaml_files_load_unchanged():
    from imaginaire_b200.config import Config
    from imaginaire_b200.generators.spade import Generator
    for rel in ('configs/unit_test/spade.yaml', 'configs/projects/spade/cocostuff/base128_bs4.yaml'):
        cfg = Config(os.path.join('/root/reference', rel))
        assert cfg.gen.type == 'imaginaire.generators.spade'
    cfg = Config('/root/reference/configs/unit_test/spade.yaml')
    net = Generator(cfg.gen, cfg.data)
    assert sum(p.numel() for p in net.parameters()) == 109247811


def _flow_generator_cfgs():
    import types
    ns = types.SimpleNamespace
    data_cfg = ns(input_types=[{'images': ns(num_channels=3)}, {'seg_maps': ns(num_channels=12)}],
                  input_image=['images'], input_labels=['seg_maps'], num_frames_G=3)
    flow_cfg = ns(num_filters=16, max_num_filters=64, num_downsamples=2, kernel_size=3, num_res_blocks=2,
                  flow_output_multiplier=20, activation_norm_type='sync_batch', weight_norm_type='spectral')
    return flow_cfg, data_cfg


def test_flow_generator_seeded_build_reproduces_reference_weights():
    """vid2vid FlowGenerator (ref: imaginaire/generators/vid2vid.py:390-481): same state-dict keys in
    the same order and, built with the same seed + the trainer's weights_init, bit-identical
    parameters to the reference module (golden made by oracle/make_golden_vid2vid.py)."""
    from imaginaire_b200.generators.vid2vid import FlowGenerator
    from imaginaire_b200.utils.init_weight import weights_init
    d = H.load_golden('flow_generator_a.npz')
    torch.manual_seed(11)
    net = FlowGenerator(*_flow_generator_cfgs())
    net.apply(weights_init('xavier', 0.02))
    sd = net.state_dict()
    assert list(sd.keys()) == [str(k) for k in d['keys']]
    ref = H.split_sd(d, 'sd.')
    for k in sd:
        assert torch.equal(sd[k], ref[k]), k
    net.load_state_dict(ref, strict=True)


def test_host_side_planners_and_validation_of_the_later_entry_points():
    """Entry points that decide or validate on the host can be exercised without a GPU."""
    import ctypes
    from imaginaire_b200 import _C
    lib = _C.lib()
    # correlation geometry (ref: correlation_cuda.cc:25-35): FlowNetC's cost volume
    oc, oh, ow = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    assert lib.raw('imx_correlation_output_size')(256, 48, 64, 20, 1, 20, 1, 2, ctypes.byref(oc), ctypes.byref(oh),
                                                  ctypes.byref(ow)) == 0
    assert (oc.value, oh.value, ow.value) == (441, 48, 64)
    # a 3x3 patch with pad == max_displacement would read before the padded plane (the reference does)
    assert lib.raw('imx_correlation_output_size')(8, 8, 8, 4, 3, 4, 1, 2, ctypes.byref(oc), ctypes.byref(oh),
                                                  ctypes.byref(ow)) != 0
    assert 'outside the padded input' in lib.last_error()
    # split-K planner: a 16x32 layer with 2048 -> 128 channels wants a workspace, a 128x256 layer does not
    f = lib.raw('imx_conv2d_fprop_workspace_bytes')
    assert f(8, 16, 32, 2048, 128, 128, 5, 5, 1, 2, 0) == 8 * 16 * 32 * 128 * 4
    assert f(8, 128, 256, 512, 512, 512, 3, 3, 1, 1, 0) == 0
    assert f(8, 16, 32, 2048, 128, 128, 5, 5, 1, 2, 32) == 0          # statistics epilogue: never split
    assert f(8, 16, 32, 2047, 128, 128, 5, 5, 1, 2, 0) < 0            # Cin_p must be a multiple of 8
    # optimizer: nothing to do is fine, nonsense hyper-parameters are not
    assert lib.raw('imx_adam_multi_tensor_f32')(None, None, None, None, None, 0, 1e-3, 0.9, 0.999, 1e-8, 0.0, 1, None) == 0
    one = (ctypes.c_void_p * 1)(8)
    n = (ctypes.c_int64 * 1)(4)
    assert lib.raw('imx_adam_multi_tensor_f32')(one, one, one, one, n, 1, 1e-3, 1.5, 0.999, 1e-8, 0.0, 1, None) != 0
    assert lib.raw('imx_adam_multi_tensor_f32')(one, one, one, one, n, 1, 1e-3, 0.9, 0.999, 1e-8, 0.0, 0, None) != 0
    # small-channel helpers
    assert lib.raw('imx_shift_add_taps_f32')(8, None, 8, 1, 8, 8, 80, 5, 5, 1, 3, None) != 0
    assert "'same' convolution" in lib.last_error()
    assert lib.raw('imx_shift_add_taps_f32')(8, None, 8, 1, 8, 8, 80, 5, 5, 2, 9, None) != 0
    # peer-memory gradient path
    assert lib.raw('imx_wgrad_reduce_peers_f32')(one, 9, 8, 0, 1, 1, 1, 1, 1, 128, 64, 1.0, 0, None) != 0
    assert lib.raw('imx_gather_rows_peers_f32')(one, 1, 0, None, 0, 1, None) == 0


def test_loss_and_pooling_helpers_have_no_cpu_path():
    """The fused loss / pooling / resize helpers keep torch expressions only for CUDA tensors the
    kernels do not cover; CPU tensors are refused like everywhere else in the product path."""
    import torch
    from imaginaire_b200 import functional as NF
    calls = ((NF.l1_loss_mean, (torch.zeros(2), torch.zeros(2))),
             (NF.max_pool2x2, (torch.zeros(1, 8, 2, 2),)),
             (NF.bilinear_half_align_corners, (torch.zeros(1, 1, 4, 4),)),
             (NF.channel_dot, (torch.zeros(1, 8, 2, 2), torch.zeros(1, 8, 2, 2))))
    for fn, args in calls:
        with pytest.raises(RuntimeError, match='no CPU path'):
            fn(*args)



def test_imaginaire_alias_package_and_entry_points():
    """`import imaginaire.x.y` returns the module object imaginaire_b200.x.y, so the reference's config files
    (cfg.gen.type / cfg.dis.type / cfg.trainer.type name modules by path, train.py:33-60) and user code resolve to the
    native modules; the trainer and the factory functions have the reference's signatures
    (trainers/base.py:42-51, utils/trainer.py:40-125)."""
    import importlib
    import inspect
    import imaginaire
    import imaginaire_b200.generators.spade as g
    import imaginaire_b200.layers as layers
    assert importlib.import_module('imaginaire.generators.spade') is g
    assert importlib.import_module('imaginaire.layers') is layers
    from imaginaire.layers import Conv2dBlock, Res2dBlock
    assert Conv2dBlock is layers.Conv2dBlock and Res2dBlock is layers.Res2dBlock
    from imaginaire.third_party.resample2d.resample2d import Resample2d
    import imaginaire_b200.third_party.resample2d.resample2d as r2d
    assert Resample2d is r2d.Resample2d
    with pytest.raises(ModuleNotFoundError):
        importlib.import_module('imaginaire.does_not_exist')
    from imaginaire_b200.config import Config
    cfg = Config(os.path.join(ROOT, 'configs', 'spade_c3_256x512.yaml'))
    for key in (cfg.gen.type, cfg.dis.type, cfg.trainer.type):
        assert importlib.import_module(key).__name__.startswith('imaginaire_b200.')
    trainer_cls = importlib.import_module(cfg.trainer.type).Trainer
    assert list(inspect.signature(trainer_cls.__init__).parameters)[1:] == \
        ['cfg', 'net_G', 'net_D', 'opt_G', 'opt_D', 'sch_G', 'sch_D', 'train_data_loader', 'val_data_loader']
    from imaginaire.utils import trainer as ut
    assert list(inspect.signature(ut.get_trainer).parameters) == \
        ['cfg', 'net_G', 'net_D', 'opt_G', 'opt_D', 'sch_G', 'sch_D', 'train_data_loader', 'val_data_loader']
    for name in ('gen_update', 'dis_update', 'start_of_iteration', 'end_of_iteration', 'start_of_epoch', 'end_of_epoch',
                 'save_checkpoint', 'load_checkpoint', 'gen_forward', 'dis_forward'):
        assert callable(getattr(trainer_cls, name)), name


def test_schedulers_and_checkpoint_round_trip(tmp_path):
    """get_scheduler (utils/trainer.py:219-240) and the checkpoint format of trainers/base.py:199-265, 795-829 on a
    toy pair of networks (CPU: no kernel is launched)."""
    import types
    import torch
    from imaginaire_b200.config import AttrDict
    from imaginaire_b200.trainers.base import BaseTrainer, _save_checkpoint
    from imaginaire_b200.utils.trainer import WrappedModel, get_scheduler

    class Toy(BaseTrainer):
        def _init_loss(self, cfg):
            pass
    cfg = AttrDict(trainer=AttrDict(model_average=False), logdir=str(tmp_path), snapshot_save_iter=10 ** 9,
                   snapshot_save_epoch=10 ** 9, logging_iter=100,
                   gen_opt=AttrDict(lr_policy=AttrDict(type='step', step_size=2, gamma=0.5, iteration_mode=True)),
                   dis_opt=AttrDict(lr_policy=AttrDict(type='constant', iteration_mode=False)))
    net_G, net_D = WrappedModel(torch.nn.Linear(3, 2)), WrappedModel(torch.nn.Linear(2, 1))
    opt_G, opt_D = torch.optim.Adam(net_G.parameters(), lr=1e-2), torch.optim.Adam(net_D.parameters(), lr=1e-2)
    sch_G, sch_D = get_scheduler(cfg.gen_opt, opt_G), get_scheduler(cfg.dis_opt, opt_D)
    t = Toy(cfg, net_G, net_D, opt_G, opt_D, sch_G, sch_D, (), None)
    t.start_iteration_time = 0.0
    for it in range(1, 4):
        opt_G.step()
        t.end_of_iteration({}, 0, it)
    assert abs(sch_G.get_last_lr()[0] - 0.5e-2) < 1e-12 and sch_D.get_last_lr()[0] == 1e-2
    path = t.save_checkpoint(3, 7)
    assert os.path.basename(path) == 'epoch_00003_iteration_000000007_checkpoint.pt'
    assert open(os.path.join(str(tmp_path), 'latest_checkpoint.txt')).read() == 'latest_checkpoint: ' + os.path.basename(path)
    ck = torch.load(path, weights_only=False)
    assert set(ck) == {'net_G', 'net_D', 'opt_G', 'opt_D', 'sch_G', 'sch_D', 'current_epoch', 'current_iteration'}
    assert list(ck['net_G']) == ['module.weight', 'module.bias']           # the wrapper's prefix, as in the reference
    with torch.no_grad():
        net_G.module.weight.zero_()
    assert t.load_checkpoint(cfg, 'no/such/file') == (3, 7)                # resumes from latest_checkpoint.txt
    assert torch.equal(net_G.module.weight, ck['net_G']['module.weight'])
