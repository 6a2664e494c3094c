"""Training loss of GraspNet (reference: models/loss.py:22-131, utils/loss_utils.py:98-114), restated so that the
training step of BASELINE.json config 5 runs without the reference tree: objectness cross-entropy + view-score MSE +
0.2 x (grasp score Huber + in-plane angle cross-entropy + width Huber + tolerance Huber).  Same end_points keys in
and out ('loss/...' and the accuracy statistics)."""
import torch
import torch.nn.functional as F

GRASP_MAX_WIDTH = 0.1            # utils/loss_utils.py:15-18
GRASP_MAX_TOLERANCE = 0.05
THRESH_GOOD = 0.7
THRESH_BAD = 0.1


def huber_loss(error, delta=1.0):
    """0.5 x^2 inside |x| <= delta, linear outside (loss_utils.py:98-114)"""
    a = error.abs()
    q = a.clamp(max=delta)
    return 0.5 * q ** 2 + delta * (a - q)


def _seed_objectness(end_points):
    """(B, Ns) objectness label of the seed points (the per-point labels gathered at the seeds' indices)"""
    return torch.gather(end_points['objectness_label'], 1, end_points['fp2_inds'].long())


def compute_objectness_loss(end_points):
    score, label = end_points['objectness_score'], _seed_objectness(end_points)
    loss = F.cross_entropy(score, label, reduction='mean')
    pred = score.argmax(1)
    hit = pred == label.long()
    end_points['loss/stage1_objectness_loss'] = loss
    end_points['stage1_objectness_acc'] = hit.float().mean()
    end_points['stage1_objectness_prec'] = hit[pred == 1].float().mean()
    end_points['stage1_objectness_recall'] = hit[label == 1].float().mean()
    return loss, end_points


def compute_view_loss(end_points):
    view_score, view_label = end_points['view_score'], end_points['batch_grasp_view_label']
    on_object = (_seed_objectness(end_points) > 0).unsqueeze(-1).expand_as(view_label)
    loss = F.mse_loss(view_score, view_label, reduction='none')[on_object].mean()
    end_points['loss/stage1_view_loss'] = loss
    end_points['stage1_pos_view_pred_count'] = ((view_score >= THRESH_GOOD) & on_object).long().sum()
    return loss, end_points


def compute_grasp_loss(end_points, use_template_in_training=True):
    label = end_points['batch_grasp_label']                  # (B, Ns, A, D), of the seed's top view
    offset = end_points['batch_grasp_offset']                # (B, Ns, A, D, 3): angle, depth, width
    tolerance = end_points['batch_grasp_tolerance']          # (B, Ns, A, D)
    A = label.size(2)
    best = label.argmax(dim=2, keepdim=True)                 # (B, Ns, 1, D): the best in-plane angle per depth
    at_best = lambda t: torch.gather(t, 2, best).squeeze(2)  # (B, Ns, D)
    target_label, target_width, target_tol = at_best(label), at_best(offset[..., 2]), at_best(tolerance)
    weight = ((target_label > THRESH_BAD) & _seed_objectness(end_points).bool().unsqueeze(-1)).float()
    norm = weight.sum() + 1e-6
    masked_mean = lambda t: (t * weight).sum() / norm
    best_t = best.transpose(1, 2)                            # (B, 1, Ns, D): index into the (B, A, Ns, D) predictions
    pred_at_best = lambda key: torch.gather(end_points[key], 1, best_t).squeeze(1)

    score_loss = masked_mean(huber_loss(pred_at_best('grasp_score_pred') - target_label, delta=1.0))
    angle_cls = best.squeeze(2)
    angle_score = end_points['grasp_angle_cls_pred']
    angle_loss = masked_mean(F.cross_entropy(angle_score, angle_cls, reduction='none'))
    width_loss = masked_mean(huber_loss((pred_at_best('grasp_width_pred') - target_width) / GRASP_MAX_WIDTH, delta=1))
    tol_loss = masked_mean(huber_loss((pred_at_best('grasp_tolerance_pred') - target_tol) / GRASP_MAX_TOLERANCE, delta=1))

    end_points['loss/stage2_grasp_score_loss'] = score_loss
    end_points['loss/stage2_grasp_angle_class_loss'] = angle_loss
    end_points['loss/stage2_grasp_width_loss'] = width_loss
    end_points['loss/stage2_grasp_tolerance_loss'] = tol_loss
    diff = (angle_score.argmax(1) - angle_cls).abs()
    on = weight.bool()
    for name, k in (("0_degree", 0), ("15_degree", 1), ("30_degree", 2)):
        end_points['stage2_grasp_angle_class_acc/' + name] = ((diff <= k) | (diff >= A - k) if k else diff == 0)[on].float().mean()
    return score_loss + angle_loss + width_loss + tol_loss, end_points


def get_loss(end_points):
    objectness, end_points = compute_objectness_loss(end_points)
    view, end_points = compute_view_loss(end_points)
    grasp, end_points = compute_grasp_loss(end_points)
    loss = objectness + view + 0.2 * grasp
    end_points['loss/overall_loss'] = loss
    return loss, end_points
