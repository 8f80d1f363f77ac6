from .sgd import SGD
from .adam import Adam, FlatAdam


def build_optimizer(cfg, params):
    if cfg.optimizer == "adam":
        return Adam(params, lr=cfg.lr, weight_decay=cfg.weight_decay)
    return SGD(params, lr=cfg.lr, momentum=cfg.momentum, weight_decay=cfg.weight_decay,
               nesterov=cfg.nesterov)
