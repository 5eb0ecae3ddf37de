"""Topic-model networks as ``torch.nn`` modules.

These modules are (a) the parameter containers whose ``state_dict()`` keys and shapes are the
checkpoint / wire format of the framework, and (b) the eager PyTorch execution path used on CPU
and as the numerical oracle for the sm_100a kernels.  On a CUDA device training goes through
``gfedntm_b200.engine`` which re-homes these parameters into flat buffers and runs the fused
kernels on them; the modules keep working on the same storage.

State-dict naming follows the reference so checkpoints and ``ModelUpdate`` messages interoperate:
``prior_mean``, ``prior_variance``, ``beta``, ``inf_net.input_layer.*``, ``inf_net.hiddens.l_<i>.0.*``,
``inf_net.f_mu.*``, ``inf_net.f_mu_batchnorm.*``, ``inf_net.f_sigma.*``, ``inf_net.f_sigma_batchnorm.*``,
``beta_batchnorm.*``, ``inf_net.adapt_bert.*`` (CombinedTM), ``label_classification.*``.

Ref: src/models/base/pytorchavitm/avitm_network/{inference_network,decoder_network}.py,
src/models/base/contextualized_topic_models/ctm_network/{inference_network,decoding_network}.py.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Optional, Sequence, Tuple

import torch
from torch import nn
from torch.nn import functional as F

ACTIVATION_NAMES = ("softplus", "relu", "sigmoid", "tanh", "leakyrelu", "rrelu", "elu", "selu")
ENCODER_DROPOUT = 0.2  # the reference never forwards `dropout` to its inference networks
                       # (decoder_network.py:65-66), so the encoder dropout is always the default


def make_activation(name: str) -> nn.Module:
    table = {"softplus": nn.Softplus, "relu": nn.ReLU, "sigmoid": nn.Sigmoid, "tanh": nn.Tanh,
             "leakyrelu": nn.LeakyReLU, "rrelu": nn.RReLU, "elu": nn.ELU, "selu": nn.SELU}
    if name not in table:
        raise ValueError(f"activation must be one of {ACTIVATION_NAMES}, got {name!r}")
    return table[name]()


class InferenceNetwork(nn.Module):
    """Encoder q(theta | doc).  ``kind`` selects the first layer:

    * ``"bow"``      — ProdLDA / NeuralLDA: ``input_layer = Linear(V, H0)`` on the BoW vector.
    * ``"zeroshot"`` — ZeroShotTM: ``input_layer = Linear(C + L, H0)`` on the contextual embedding.
    * ``"combined"`` — CombinedTM: ``adapt_bert = Linear(C, V)``; ``input_layer = Linear(2V + L, H0)``
      on ``cat(bow, adapt_bert(embedding)[, labels])`` (no non-linearity in between).
    """

    def __init__(self, input_size: int, output_size: int, hidden_sizes: Sequence[int],
                 activation: str = "softplus", dropout: float = ENCODER_DROPOUT, kind: str = "bow",
                 contextual_size: int = 0, label_size: int = 0) -> None:
        super().__init__()
        if kind not in ("bow", "zeroshot", "combined"):
            raise ValueError("kind must be 'bow', 'zeroshot' or 'combined'")
        hidden_sizes = tuple(int(h) for h in hidden_sizes)
        if len(hidden_sizes) < 1:
            raise ValueError("hidden_sizes needs at least one layer")
        if dropout < 0:
            raise ValueError("dropout must be >= 0")
        self.kind = kind
        self.input_size = int(input_size)
        self.output_size = int(output_size)
        self.hidden_sizes = hidden_sizes
        self.contextual_size = int(contextual_size)
        self.label_size = int(label_size)
        self.dropout = float(dropout)
        self.activation_name = activation
        self.activation = make_activation(activation)

        if kind == "bow":
            first_in = self.input_size
        elif kind == "zeroshot":
            first_in = self.contextual_size + self.label_size
        else:
            self.adapt_bert = nn.Linear(self.contextual_size, self.input_size)
            first_in = 2 * self.input_size + self.label_size
        self.input_layer = nn.Linear(first_in, hidden_sizes[0])
        self.hiddens = nn.Sequential(OrderedDict(
            (f"l_{i}", nn.Sequential(nn.Linear(h_in, h_out), self.activation))
            for i, (h_in, h_out) in enumerate(zip(hidden_sizes[:-1], hidden_sizes[1:]))))
        self.f_mu = nn.Linear(hidden_sizes[-1], self.output_size)
        self.f_mu_batchnorm = nn.BatchNorm1d(self.output_size, affine=False)
        self.f_sigma = nn.Linear(hidden_sizes[-1], self.output_size)
        self.f_sigma_batchnorm = nn.BatchNorm1d(self.output_size, affine=False)
        self.dropout_enc = nn.Dropout(p=self.dropout)

    def first_layer_input(self, x_bow: Optional[torch.Tensor], x_ctx: Optional[torch.Tensor],
                          labels: Optional[torch.Tensor]) -> torch.Tensor:
        if self.kind == "bow":
            return x_bow
        if self.kind == "zeroshot":
            return x_ctx if labels is None else torch.cat((x_ctx, labels), 1)
        parts = [x_bow, self.adapt_bert(x_ctx)]
        if labels is not None:
            parts.append(labels)
        return torch.cat(parts, 1)

    def forward(self, x_bow=None, x_ctx=None, labels=None) -> Tuple[torch.Tensor, torch.Tensor]:
        h = self.activation(self.input_layer(self.first_layer_input(x_bow, x_ctx, labels)))
        h = self.hiddens(h)
        h = self.dropout_enc(h)
        mu = self.f_mu_batchnorm(self.f_mu(h))
        log_sigma = self.f_sigma_batchnorm(self.f_sigma(h))
        return mu, log_sigma


class DecoderNetwork(nn.Module):
    """VAE topic model: encoder + reparameterised softmax(theta) + (Prod)LDA decoder.

    ``forward`` returns ``(prior_mean, prior_variance, posterior_mean, posterior_variance,
    posterior_log_variance, word_dist[, estimated_labels])`` like the reference modules.
    """

    def __init__(self, input_size: int, n_components: int = 10, model_type: str = "prodLDA",
                 hidden_sizes: Sequence[int] = (100, 100), activation: str = "softplus",
                 dropout: float = 0.2, learn_priors: bool = True, topic_prior_mean: float = 0.0,
                 topic_prior_variance: Optional[float] = None, infnet: str = "bow",
                 contextual_size: int = 0, label_size: int = 0) -> None:
        super().__init__()
        if not (isinstance(input_size, int) and input_size > 0):
            raise ValueError("input_size must be an int > 0")
        if not (isinstance(n_components, int) and n_components > 0):
            raise ValueError("n_components must be an int > 0")
        if model_type.lower() not in ("lda", "prodlda"):
            raise ValueError("model_type must be 'prodLDA' or 'LDA'")
        if not isinstance(hidden_sizes, tuple):
            raise ValueError("hidden_sizes must be a tuple")
        if activation not in ACTIVATION_NAMES:
            raise ValueError(f"activation must be one of {ACTIVATION_NAMES}")
        if dropout < 0:
            raise ValueError("dropout must be >= 0")

        self.input_size = input_size
        self.n_components = n_components
        self.model_type = model_type
        self.hidden_sizes = tuple(hidden_sizes)
        self.activation = activation
        self.dropout = dropout
        self.learn_priors = learn_priors
        self.infnet = infnet
        self.contextual_size = contextual_size
        self.label_size = label_size
        # plain attribute (not a registered parameter): assigning the nn.Parameter through
        # Module.__setattr__ would add a 'topic_word_matrix' key to the state dict
        object.__setattr__(self, "topic_word_matrix", None)

        self.inf_net = InferenceNetwork(input_size, n_components, hidden_sizes, activation,
                                        kind=infnet, contextual_size=contextual_size,
                                        label_size=label_size)
        if label_size != 0:
            self.label_classification = nn.Linear(n_components, label_size)

        # Laplace approximation of a symmetric Dirichlet(1) prior in the softmax basis
        if topic_prior_variance is None:
            topic_prior_variance = 1.0 - 1.0 / n_components
        pm = torch.full((n_components,), float(topic_prior_mean))
        pv = torch.full((n_components,), float(topic_prior_variance))
        if learn_priors:
            self.prior_mean = nn.Parameter(pm)
            self.prior_variance = nn.Parameter(pv)
        else:
            self.register_buffer("prior_mean", pm, persistent=False)
            self.register_buffer("prior_variance", pv, persistent=False)

        self.beta = nn.Parameter(torch.empty(n_components, input_size))
        nn.init.xavier_uniform_(self.beta)
        self.beta_batchnorm = nn.BatchNorm1d(input_size, affine=False)
        self.drop_theta = nn.Dropout(p=dropout)

    # -- pieces -------------------------------------------------------------------------------
    @staticmethod
    def reparameterize(mu: torch.Tensor, logvar: torch.Tensor, eps: Optional[torch.Tensor] = None):
        std = torch.exp(0.5 * logvar)
        if eps is None:
            eps = torch.randn_like(std)
        return mu + eps * std

    def decode(self, theta: torch.Tensor) -> torch.Tensor:
        if self.model_type.lower() == "prodlda":
            word_dist = F.softmax(self.beta_batchnorm(theta @ self.beta), dim=1)
            object.__setattr__(self, "topic_word_matrix", self.beta)
        else:
            beta = F.softmax(self.beta_batchnorm(self.beta), dim=1)
            word_dist = theta @ beta
            object.__setattr__(self, "topic_word_matrix", beta)
        return word_dist

    def forward(self, x_bow=None, x_ctx=None, labels=None, eps: Optional[torch.Tensor] = None):
        mu, logvar = self.inf_net(x_bow, x_ctx, labels)
        var = torch.exp(logvar)
        theta = F.softmax(self.reparameterize(mu, logvar, eps), dim=1)
        theta = self.drop_theta(theta)
        word_dist = self.decode(theta)
        out = (self.prior_mean, self.prior_variance, mu, var, logvar, word_dist)
        if self.infnet != "bow":
            est = self.label_classification(theta) if (labels is not None and self.label_size) else None
            return out + (est,)
        return out

    def get_theta(self, x_bow=None, x_ctx=None, labels=None) -> torch.Tensor:
        """One stochastic draw of theta (no theta-dropout), as the reference's ``get_theta``."""
        with torch.no_grad():
            mu, logvar = self.inf_net(x_bow, x_ctx, labels)
            return F.softmax(self.reparameterize(mu, logvar), dim=1)


def elbo_terms(inputs: torch.Tensor, word_dists: torch.Tensor, prior_mean: torch.Tensor,
               prior_variance: torch.Tensor, posterior_mean: torch.Tensor,
               posterior_variance: torch.Tensor, posterior_log_variance: torch.Tensor
               ) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-document (KL, RL).  Ref: avitm.py:203-229 / ctm.py:182-238."""
    n_components = posterior_mean.shape[1]
    var_division = torch.sum(posterior_variance / prior_variance, dim=1)
    diff = prior_mean - posterior_mean
    diff_term = torch.sum(diff * diff / prior_variance, dim=1)
    logvar_det = prior_variance.log().sum() - posterior_log_variance.sum(dim=1)
    kl = 0.5 * (var_division + diff_term - n_components + logvar_det)
    rl = -torch.sum(inputs * torch.log(word_dists + 1e-10), dim=1)
    return kl, rl
