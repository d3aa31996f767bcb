import torch
import torchvision

from pytorch_distributed_learning_b200.models import ConvNet, resnet18, resnet50


def test_convnet_shapes_and_keys():
    m = ConvNet()
    assert list(m.state_dict().keys()) == ["conv1.weight", "conv1.bias", "conv2.weight", "conv2.bias", "conv3.weight", "conv3.bias",
                                           "fc1.weight", "fc1.bias"]
    assert m(torch.randn(3, 1, 28, 28)).shape == (3, 10)
    assert len(list(m.buffers())) == 0


def test_convnet_matches_reference_definition_on_cpu():
    import torch.nn as nn

    class Ref(nn.Module):  # layer list of launch_dist.py:9-41, written from its spec
        def __init__(self):
            super().__init__()
            self.conv1, self.conv2, self.conv3 = nn.Conv2d(1, 32, 5, 1, 1), nn.Conv2d(32, 64, 3), nn.Conv2d(64, 128, 3)
            self.fc1 = nn.Linear(2048, 10)

        def forward(self, x):
            x = nn.functional.max_pool2d(torch.relu(self.conv1(x)), 2, 2)
            x = nn.functional.max_pool2d(torch.relu(self.conv2(x)), 2, 1)
            x = nn.functional.max_pool2d(torch.relu(self.conv3(x)), 2, 2)
            return self.fc1(x.view(-1, 2048))

    ours, ref = ConvNet(), Ref()
    ref.load_state_dict(ours.state_dict())
    x = torch.randn(5, 1, 28, 28)
    torch.testing.assert_close(ours(x), ref(x))


def test_resnet_state_dict_matches_torchvision():
    for ours, ref in ((resnet18(num_classes=10), torchvision.models.resnet18(num_classes=10)), (resnet50(), torchvision.models.resnet50())):
        a, b = ours.state_dict(), ref.state_dict()
        assert list(a.keys()) == list(b.keys())
        assert all(a[k].shape == b[k].shape for k in a)
    ours, ref = resnet18(num_classes=10), torchvision.models.resnet18(num_classes=10)
    ref.load_state_dict(ours.state_dict())
    x = torch.randn(2, 3, 32, 32)
    ours.eval(), ref.eval()
    torch.testing.assert_close(ours(x), ref(x))
