"""Parameter and activation memory of an arbitrary ``nn.Module`` measured with forward hooks.

Same helpers as the reference tool (``get_model_memory``, ``ActivationCounter``,
``activation_counter_hook``, ``register_hooks_recursive``) and the same demonstration: a four-layer
MLP at batch 4 prints ``Model Memory: 12,597,248 bytes`` and ``Activation Memory: 65,536 bytes``.
Dropout layers are charged one byte per element (only the mask has to be kept); the fused kernels
of this repository keep no mask at all (it is recomputed from a seed), which
``ActivationCounter(dropout_bytes_per_element=0)`` models.
"""
import torch


def get_model_memory(model: torch.nn.Module) -> int:
    """Bytes held by the parameters of `model` (shared parameters counted once)."""
    seen, total = set(), 0
    for p in model.parameters():
        if id(p) not in seen:
            seen.add(id(p))
            total += p.numel() * p.element_size()
    return total


class ActivationCounter:
    def __init__(self, dropout_bytes_per_element: int = 1):
        self.activation_bytes = 0
        self.dropout_bytes_per_element = dropout_bytes_per_element
        self.per_module = []

    def add_activations(self, tensor):
        self.activation_bytes += tensor.numel() * tensor.element_size()

    def add_activation_bytes(self, nbytes):
        self.activation_bytes += nbytes


def activation_counter_hook(counter: ActivationCounter):
    def hook(module, _inputs, output):
        out = output[0] if isinstance(output, (tuple, list)) else output
        if type(module).__name__ == "Dropout":
            nbytes = out.numel() * counter.dropout_bytes_per_element
            counter.add_activation_bytes(nbytes)
        else:
            nbytes = out.numel() * out.element_size()
            counter.add_activations(out)
        counter.per_module.append((type(module).__name__, nbytes))
    return hook


def register_hooks_recursive(model: torch.nn.Module, counter: ActivationCounter):
    """Hooks every descendant (children, grandchildren, ...) of `model`; returns the handles."""
    handles = []
    for child in model.children():
        handles.append(child.register_forward_hook(activation_counter_hook(counter)))
        handles.extend(register_hooks_recursive(child, counter))
    return handles


def main():
    model = torch.nn.Sequential(torch.nn.Linear(512, 1024), torch.nn.Linear(1024, 1024),
                                torch.nn.Linear(1024, 1024), torch.nn.Linear(1024, 512))
    print("Model Memory: {:,} bytes".format(get_model_memory(model)))
    counter = ActivationCounter()
    register_hooks_recursive(model, counter)
    inputs = torch.randn(4, 512)
    model(inputs)
    counter.add_activations(inputs)   # hooks only see outputs: add the input tensor itself
    print("Activation Memory: {:,} bytes".format(counter.activation_bytes))


if __name__ == "__main__":
    main()
