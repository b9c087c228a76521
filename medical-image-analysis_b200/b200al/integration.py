"""Hook the drop-in selectors into the reference's trainer without editing it.

`ALTrainer._setup_active_selector` (src/training/al_trainer.py:802-883) picks a
selector by `config.active_selector_name` and raises ValueError for unknown
names.  `install_into_al_trainer(ALTrainer)` wraps that method so the names in
`b200al.SELECTORS` ("b200-entropy", "b200-coreset-l2", ...) construct the
B200 selectors with the same kwargs the reference passes (batch_size,
num_workers, pin_memory, and coreset_criteria / coreset_fusion for the coreset
names); every other name goes to the original method.
"""
from __future__ import annotations


def install_into_al_trainer(trainer_cls, reference_abc=None):
    from .selectors import SELECTORS, ActiveSelector

    original = trainer_cls._setup_active_selector

    def _setup_active_selector(self):
        name = self.config.active_selector_name
        if name in SELECTORS:
            kw = dict(
                batch_size=self.config.batch_size,
                num_workers=self.config.num_workers,
                pin_memory=self.config.pin_memory,
            )
            if name.startswith("b200-coreset"):  # al_trainer.py:823-844 forwards these two to CoresetSelector
                kw["coreset_criteria"] = getattr(self.config, "coreset_criteria", "min")
                kw["coreset_fusion"] = getattr(self.config, "coreset_fusion", "add")
                kw["feature_path"] = getattr(self.config, "feature_path", None)
                kw["loaded_feature_weight"] = getattr(self.config, "loaded_feature_weight", 0.0)
            if name.startswith("b200-kmean"):  # al_trainer.py:845-870
                kw["coreset_criteria"] = getattr(self.config, "coreset_criteria", "min")
                kw["sharp_factor"] = getattr(self.config, "kmean_sharp_factor", 1.0)
                kw["softmax"] = getattr(self.config, "kmean_softmax", False)
                kw["feature_path"] = getattr(self.config, "feature_path", None)
                kw["loaded_feature_weight"] = getattr(self.config, "loaded_feature_weight", 1.0)
                kw["loaded_feature_only"] = getattr(self.config, "loaded_feature_only", False)
            if name == "b200-badge":  # al_trainer.py:871-879
                kw.update(dice_loss=self.supervised_loss.dice_loss, ce_loss=self.supervised_loss.ce_loss,
                          multiple_loss="add")
            self.active_selector = SELECTORS[name](**kw)
        else:
            original(self)

    trainer_cls._setup_active_selector = _setup_active_selector
    if reference_abc is not None:  # make isinstance(sel, reference ActiveSelector) hold
        reference_abc.register(ActiveSelector)
    return trainer_cls
