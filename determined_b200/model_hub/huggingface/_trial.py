"""``BaseTransformerTrial``: a PyTorchTrial that builds a transformers model, optimizer and LR schedule
from hyperparameters (reference ``model_hub/model_hub/huggingface/_trial.py:1-303``).

B200 notes: the optimizer it builds is a ``torch.optim.AdamW`` so ``wrap_optimizer`` can convert it to
the fused flat-buffer kernel; ``use_apex_amp`` maps to native bf16 autocast (apex is not needed);
``model_mode: "native-gpt"`` builds :mod:`determined_b200.models.gpt` (tcgen05 GEMMs + attention) from
``gpt_preset`` instead of a transformers Auto class.
"""
import logging
from typing import Any, Dict, List, Optional, Tuple, Union

import torch

import determined_b200.pytorch as det_torch
from determined_b200.model_hub import utils
from determined_b200.model_hub.huggingface import _config_parser as hf_parse

logger = logging.getLogger("determined.model_hub")


def _model_modes() -> Dict[str, Any]:
    import transformers

    return {"base": transformers.AutoModel, "pretraining": transformers.AutoModelForPreTraining, "causal-lm": transformers.AutoModelForCausalLM,
            "masked-lm": transformers.AutoModelForMaskedLM, "seq2seq-lm": transformers.AutoModelForSeq2SeqLM,
            "sequence-classification": transformers.AutoModelForSequenceClassification, "multiple-choice": transformers.AutoModelForMultipleChoice,
            "next-sentence": transformers.AutoModelForNextSentencePrediction, "token-classification": transformers.AutoModelForTokenClassification,
            "question-answering": transformers.AutoModelForQuestionAnswering}


def build_using_auto(config_kwargs: Dict, tokenizer_kwargs: Dict, model_mode: str, model_kwargs: Dict, use_pretrained_weights: bool = True) -> Tuple[Any, Any, Any]:
    import transformers

    config = transformers.AutoConfig.from_pretrained(**config_kwargs)
    tokenizer = transformers.AutoTokenizer.from_pretrained(**tokenizer_kwargs)
    builder = _model_modes()[model_mode]
    if use_pretrained_weights:
        model = builder.from_pretrained(**dict(model_kwargs, config=config))
    else:
        model = builder.from_config(config)
    return config, tokenizer, model


def group_parameters_for_optimizer(model: torch.nn.Module, weight_decay: Optional[float] = 0, no_decay: Tuple[str, ...] = ("bias", "LayerNorm.weight", "ln")) -> List[Dict[str, Any]]:
    decay, nodecay = [], []
    for n, p in model.named_parameters():
        if not p.requires_grad:
            continue
        (nodecay if any(nd in n for nd in no_decay) else decay).append(p)
    return [{"params": decay, "weight_decay": weight_decay}, {"params": nodecay, "weight_decay": 0.0}]


def build_default_optimizer(model: torch.nn.Module, optimizer_kwargs: hf_parse.OptimizerKwargs) -> torch.optim.Optimizer:
    groups = group_parameters_for_optimizer(model, optimizer_kwargs.weight_decay)
    if optimizer_kwargs.adafactor:
        import transformers.optimization as hf_opt

        return hf_opt.Adafactor(groups, lr=optimizer_kwargs.learning_rate, scale_parameter=optimizer_kwargs.scale_parameter, relative_step=optimizer_kwargs.relative_step)
    return torch.optim.AdamW(groups, lr=optimizer_kwargs.learning_rate, betas=(optimizer_kwargs.adam_beta1, optimizer_kwargs.adam_beta2), eps=optimizer_kwargs.adam_epsilon)


def build_default_lr_scheduler(optimizer: torch.optim.Optimizer, scheduler_kwargs: hf_parse.LRSchedulerKwargs) -> Any:
    try:
        import transformers.optimization as hf_opt

        return hf_opt.get_scheduler(scheduler_kwargs.lr_scheduler_type, optimizer, num_warmup_steps=scheduler_kwargs.num_warmup_steps,
                                    num_training_steps=scheduler_kwargs.num_training_steps)
    except ImportError:
        w, n = scheduler_kwargs.num_warmup_steps, scheduler_kwargs.num_training_steps

        def linear(step: int) -> float:
            return step / max(1, w) if step < w else max(0.0, (n - step) / max(1, n - w))

        return torch.optim.lr_scheduler.LambdaLR(optimizer, linear)


class BaseTransformerTrial(det_torch.PyTorchTrial):
    """Fill in ``build_training_data_loader`` / ``build_validation_data_loader`` / ``evaluate_batch`` in a subclass."""

    def __init__(self, context: det_torch.PyTorchTrialContext) -> None:
        self.context = context
        if not hasattr(self, "hparams"):
            self.hparams = utils.AttrDict(context.get_hparams())
        if not hasattr(self, "data_config"):
            self.data_config = utils.AttrDict(context.get_data_config())
        if not hasattr(self, "exp_config"):
            self.exp_config = utils.AttrDict(context.get_experiment_config())
        self.check_hparams()
        optimizer_kwargs, scheduler_kwargs = hf_parse.default_parse_optimizer_lr_scheduler_kwargs(self.hparams)
        if self.hparams.model_mode == "native-gpt":
            from determined_b200.models import gpt

            self.config, self.tokenizer = gpt.config_for(self.hparams.get("gpt_preset", "gpt2"), **dict(self.hparams.get("gpt_overrides", {}))), None
            self.model = gpt.GPT(self.config)
        else:
            self.config_kwargs, self.tokenizer_kwargs, self.model_kwargs = hf_parse.default_parse_config_tokenizer_model_kwargs(self.hparams)
            self.config, self.tokenizer, self.model = build_using_auto(self.config_kwargs, self.tokenizer_kwargs, self.hparams.model_mode, self.model_kwargs,
                                                                       use_pretrained_weights=self.hparams.use_pretrained_weights)
        self.model = self.context.wrap_model(self.model)
        self.optimizer = self.context.wrap_optimizer(build_default_optimizer(self.model, optimizer_kwargs))
        if self.hparams.use_apex_amp:
            self.model, self.optimizer = self.context.configure_apex_amp(models=self.model, optimizers=self.optimizer)
        self.lr_scheduler = self.context.wrap_lr_scheduler(build_default_lr_scheduler(self.optimizer, scheduler_kwargs),
                                                           det_torch.LRScheduler.StepMode.STEP_EVERY_BATCH)
        self.grad_clip_fn = None
        if optimizer_kwargs.max_grad_norm > 0:
            self.grad_clip_fn = det_torch.FusedClipGradNorm(optimizer_kwargs.max_grad_norm) if hasattr(det_torch, "FusedClipGradNorm") else \
                (lambda params: torch.nn.utils.clip_grad_norm_(params, optimizer_kwargs.max_grad_norm))

    def check_hparams(self) -> None:
        if not isinstance(self.hparams, utils.AttrDict):
            self.hparams = utils.AttrDict(self.hparams)
        if "num_training_steps" not in self.hparams:
            self.hparams.num_training_steps = utils.compute_num_training_steps(self.context.get_experiment_config(), self.context.get_global_batch_size())
        if "use_pretrained_weights" not in self.hparams:
            logger.warning("Using pretrained weights by default; set use_pretrained_weights: false to train from scratch.")
            self.hparams.use_pretrained_weights = True
        for hp in ("use_apex_amp", "model_mode", "num_training_steps"):
            assert hp in self.hparams, f"{hp} is a required hyperparameter for BaseTransformerTrial"

    def train_batch(self, batch: Any, epoch_idx: int, batch_idx: int) -> Union[torch.Tensor, Dict[str, Any]]:
        if self.hparams.model_mode == "native-gpt":
            ids = batch["input_ids"] if isinstance(batch, dict) else batch
            _, loss = self.model(ids[:, :-1], ids[:, 1:])
        else:
            outputs = self.model(**batch)
            loss = outputs["loss"] if isinstance(outputs, dict) else outputs[0]
        self.context.backward(loss)
        self.context.step_optimizer(self.optimizer, self.grad_clip_fn)
        return {"loss": loss}


def default_load_dataset(data_config_input: Any) -> Any:
    """See :func:`determined_b200.model_hub.huggingface._utils.default_load_dataset` (kept importable from here, where the
    reference defines it)."""
    from determined_b200.model_hub.huggingface import _utils

    return _utils.default_load_dataset(data_config_input)
