"""Host-side mirror of the two loss helpers the path uses (DexGanGrasp/models/losses.py:5-16,61-76)."""
import torch

from . import utils


def accuracy_evaluator(pred_label, gt_label):
    """losses.py:5-16"""
    correct = torch.eq(pred_label, gt_label)
    true_positive = torch.sum(correct * gt_label) / torch.sum(gt_label)
    true_negative = torch.sum(correct * (1. - gt_label)) / torch.sum(1. - gt_label)
    acc = correct.sum() / float(len(gt_label))
    return true_positive, true_negative, acc


def transl_rot_6D_l2_loss(pred_transl_rot_6D, gt_transl_rot_matrix, torch_l2_loss_fn, device='cpu'):
    """losses.py:61-76"""
    pred_rot_matrix = utils.rot_matrix_from_ortho6d(pred_transl_rot_6D['rot_6D'])
    pred_rot_matrix = pred_rot_matrix.view(pred_rot_matrix.shape[0], -1)
    gt_rot_matrix = gt_transl_rot_matrix['rot_matrix']
    l2_loss_rot = torch_l2_loss_fn(pred_rot_matrix, gt_rot_matrix)
    l2_loss_transl = torch_l2_loss_fn(pred_transl_rot_6D['transl'], gt_transl_rot_matrix['transl'])
    return l2_loss_transl, l2_loss_rot
