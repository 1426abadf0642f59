"""Importance-sampling pass (``training.n_importance > 0``; dormant in every reference config)."""


def render_with_importance(model, rays_o, rays_d, target_rgb, target_d, with_losses, noise):
    raise NotImplementedError('training.n_importance > 0 is not wired up yet')
