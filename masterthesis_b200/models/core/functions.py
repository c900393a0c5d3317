"""Name -> layer factories, weight init, device placement and LR schedulers.

Mirrors the construction API of the reference's src/models/core/functions.py:11-121 (same function
names, arguments and error behaviour).  The returned "layers" are structural descriptors: they keep
the reference's module tree and state_dict keys, while the arithmetic is executed by the fused
CUDA plans of the enclosing block (masterthesis_b200/functional.py).
"""
import functools
from inspect import isclass

import torch
import torch.nn as nn
from torch.nn import init
from torch.optim import lr_scheduler

from .norm import AdaptiveInstanceNorm, LayerNorm


def get_norm_layer(norm_layer='instance'):
    """class (or partial) building the named normalisation layer; None passes through
    (reference functions.py:11-26)."""
    if isinstance(norm_layer, str):
        table = {
            'batch': functools.partial(nn.BatchNorm2d, affine=True, track_running_stats=True),
            'instance': functools.partial(nn.InstanceNorm2d, affine=False, track_running_stats=False),
            'layer': functools.partial(LayerNorm),
            'adain': functools.partial(AdaptiveInstanceNorm),
        }
        if norm_layer not in table:
            raise NotImplementedError(f"norm type '{norm_layer}' is not supported at the moment")
        return table[norm_layer]
    if norm_layer is not None and isclass(norm_layer) and not issubclass(norm_layer, nn.Module):
        raise ValueError("parameter type of norm_layer should be one of 'str' or 'nn.Module' class, "
                         f"but got {norm_layer}.")
    return norm_layer


def get_activation_layer(activation=None):
    """reference functions.py:28-43"""
    if isinstance(activation, str):
        table = {
            'relu': functools.partial(nn.ReLU, inplace=True),
            'lrelu': functools.partial(nn.LeakyReLU, inplace=True),
            'tanh': functools.partial(nn.Tanh),
            'sigmoid': functools.partial(nn.Sigmoid),
        }
        if activation not in table:
            raise NotImplementedError(f"activation type '{activation}' is not supported at the moment")
        return table[activation]
    if activation is not None and isclass(activation) and not issubclass(activation, nn.Module):
        raise ValueError("parameter type of activation should be one of 'str' or 'nn.Module', "
                         f"but got {type(activation)}.")
    return activation


def get_padding_layer(padding_type=None):
    """reference functions.py:45-58"""
    if isinstance(padding_type, str):
        table = {'reflect': functools.partial(nn.ReflectionPad2d),
                 'replicate': functools.partial(nn.ReplicationPad2d)}
        if padding_type not in table:
            raise NotImplementedError(f"padding type '{padding_type}' is not supported at the moment")
        return table[padding_type]
    if padding_type is not None and isclass(padding_type) and not issubclass(padding_type, nn.Module):
        raise ValueError("parameter type of padding_type should be one of 'str' or 'nn.Module', "
                         f"but got {type(padding_type)}.")
    return None if padding_type is None else padding_type


def act_name(module):
    """structural activation descriptor -> kernel activation name"""
    if module is None:
        return None
    if isinstance(module, nn.LeakyReLU):
        if abs(module.negative_slope - 0.01) > 1e-12:
            raise NotImplementedError("only the default LeakyReLU slope 0.01 is supported")
        return 'lrelu'
    if isinstance(module, nn.ReLU):
        return 'relu'
    if isinstance(module, nn.Tanh):
        return 'tanh'
    raise NotImplementedError(f"activation {type(module).__name__} has no CUDA epilogue")


def get_scheduler(optimizer, args, cur_it=-1):
    """reference functions.py:60-70"""
    if args.lr_policy == 'lambda':
        def rule(it):
            return 1.0 - max(0, it - args.n_iter_decay) / float(args.n_iters - args.n_iter_decay + 1)
        return lr_scheduler.LambdaLR(optimizer, lr_lambda=rule, last_epoch=cur_it)
    if args.lr_policy == 'step':
        return lr_scheduler.StepLR(optimizer, step_size=args.n_iter_decay, gamma=0.1, last_epoch=cur_it)
    raise NotImplementedError(f'Learning rate policy {args.lr_policy} is not implemented')


def init_weights(net, init_type='normal', init_gain=0.02):
    """N(0, gain) (or xavier/kaiming/orthogonal) for every module whose class name STARTS with
    'Conv', zero conv bias; Linear layers keep their default init (reference functions.py:72-94)."""
    makers = {
        'normal': lambda w: init.normal_(w, 0.0, init_gain),
        'xavier': lambda w: init.xavier_normal_(w, gain=init_gain),
        'kaiming': lambda w: init.kaiming_normal_(w, a=0, mode='fan_in'),
        'orthogonal': lambda w: init.orthogonal_(w, gain=init_gain),
    }

    def visit(m):
        name = type(m).__name__
        if (hasattr(m, 'weight') or getattr(m, 'spectral', False)) and name.startswith('Conv'):
            if init_type not in makers:
                raise NotImplementedError('initialization method [%s] is not implemented' % init_type)
            if hasattr(m, 'weight'):
                makers[init_type](m.weight.data)
            # under spectral norm the reference initialises `m.weight`, a plain tensor copy that the next
            # forward overwrites: `weight_orig` keeps torch's default init, only the bias is reset
            if getattr(m, 'bias', None) is not None:
                init.constant_(m.bias.data, 0.0)
        elif 'BatchNorm2d' in name:
            init.normal_(m.weight.data, 1.0, init_gain)
            init.constant_(m.bias.data, 0.0)

    print('initialize network with %s' % init_type)
    net.apply(visit)


def init_net(net, init_type='normal', init_gain=0.02, device='cpu', gpu_ids=[]):
    """Place a network on its device and initialise it (reference functions.py:96-106).
    The reference wraps in nn.DataParallel here; this build is one process per GPU, so the net is
    moved to the current CUDA device and data parallelism is done by masterthesis_b200.parallel."""
    if 'cuda' in str(device):
        net = net.to(torch.device('cuda', torch.cuda.current_device()))
    else:
        net = net.to(device)
    if init_type:
        init_weights(net, init_type, init_gain=init_gain)
    return net


def spectral_norm(module, name='weight', n_power_iterations=1, eps=1e-12, dim=None):
    """reference functions.py:113-121 -> torch.nn.utils.spectral_norm(module, name, 1, eps, dim) with
    dim = 1 for transposed convs, 0 otherwise.  `module` is one of this package's conv parameter
    holders (blocks.Conv2d / blocks.ConvTranspose2d); the power iteration and W / sigma run in
    `module.sn_weight()` at every forward (blocks.py of this package)."""
    if name != 'weight' or n_power_iterations != 1:
        raise NotImplementedError("spectral norm is built as the reference uses it: name='weight', one power iteration")
    if dim is not None and dim != (1 if module.transposed else 0):
        raise NotImplementedError("spectral norm over a non-default dim")
    return module.enable_spectral_norm(eps)
