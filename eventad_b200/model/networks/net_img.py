"""RGB feature branch: torchvision ResNet with forward hooks and 1x1 projection convs.  Same module tree and
state-dict keys as the reference's HookModule (/root/reference/src/dagr/model/networks/net_img.py:46-134):
`module.*` (the ResNet), `feature_dconv.{i}.*`, `output_dconv.{i}.*`.  This branch stays PyTorch/cuDNN (north_star);
it only feeds the event branch."""
import torch


class HookModule(torch.nn.Module):
    def __init__(self, module, height, width, input_channels=3, feature_layers=(), output_layers=(),
                 feature_channels=None, output_channels=None):
        torch.nn.Module.__init__(self)
        self.module = module.cpu()
        if input_channels != 3:
            c1 = self.module.conv1
            self.module.conv1 = torch.nn.Conv2d(in_channels=input_channels, out_channels=c1.out_channels,
                                                kernel_size=c1.kernel_size, padding=c1.padding, bias=False)
        self.feature_layers = feature_layers
        self.output_layers = output_layers
        self.hooks, self.features, self.outputs = [], [], []
        self.register_hooks()
        self.feature_channels, self.output_channels = [], []
        self.compute_channels_with_dummy(shape=(1, input_channels, height, width))

        self.feature_dconv = torch.nn.ModuleList()
        if feature_channels is not None:
            assert len(feature_channels) == len(self.feature_channels)
            self.feature_dconv = torch.nn.ModuleList(
                [torch.nn.Conv2d(cin, cout, kernel_size=1, stride=1, padding=0)
                 for cin, cout in zip(self.feature_channels, feature_channels)])
            self.feature_channels = feature_channels
        self.output_dconv = torch.nn.ModuleList()
        if output_channels is not None:
            assert len(output_channels) == len(self.output_channels)
            self.output_dconv = torch.nn.ModuleList(
                [torch.nn.Conv2d(cin, cout, kernel_size=1, stride=1, padding=0)
                 for cin, cout in zip(self.output_channels, output_channels)])
            self.output_channels = output_channels

    def extract_layer(self, module, layer):
        return module if len(layer) == 0 else self.extract_layer(module._modules[layer[0]], layer[1:])

    def compute_channels_with_dummy(self, shape):
        with torch.no_grad():
            was = self.module.training
            self.module.eval()
            self.module.forward(torch.zeros(shape))
            self.module.train(was)
        self.feature_channels = [f.shape[1] for f in self.features]
        self.output_channels = [o.shape[1] for o in self.outputs]
        self.features, self.outputs = [], []

    def remove_hooks(self):
        for h in self.hooks:
            h.remove()

    def register_hooks(self):
        self.features, self.outputs = [], []
        for name in self.feature_layers:
            self.hooks.append(self.extract_layer(self.module, name.split(".")).register_forward_hook(
                lambda m, i, o: self.features.append(o)))
        for name in self.output_layers:
            self.hooks.append(self.extract_layer(self.module, name.split(".")).register_forward_hook(
                lambda m, i, o: self.outputs.append(o)))

    def forward(self, x):
        self.features, self.outputs = [], []
        self.module(x)
        features = self.features
        if len(self.feature_dconv) > 0:
            features = [dconv(f) for f, dconv in zip(self.features, self.feature_dconv)]
        outputs = self.outputs
        if len(self.output_dconv) > 0:
            outputs = [dconv(o) for o, dconv in zip(self.outputs, self.output_dconv)]
        return features, outputs
