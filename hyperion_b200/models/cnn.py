"""The 2-conv CNN the reference silently benchmarks under the label "Vision
Transformer" when ``vit_b_16`` is unavailable (Phase 1/baseline_performance.ipynb
c0 L41-54): Conv7x7/2(3->64)-ReLU-MaxPool3/2 - Conv3x3(64->128)-ReLU-MaxPool2 - GAP -
Linear(128->1000); 212,328 parameters.  Kept (correctly labelled) because the
published "ViT" row was measured on it."""
from __future__ import annotations

import torch.nn as nn


class FallbackCNN(nn.Module):
    def __init__(self, num_classes: int = 1000):
        super().__init__()
        self.features = nn.Sequential(
            nn.Conv2d(3, 64, kernel_size=7, stride=2, padding=3), nn.ReLU(inplace=True),
            nn.MaxPool2d(kernel_size=3, stride=2, padding=1),
            nn.Conv2d(64, 128, kernel_size=3, padding=1), nn.ReLU(inplace=True),
            nn.MaxPool2d(kernel_size=2, stride=2),
        )
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.classifier = nn.Linear(128, num_classes)

    def forward(self, x):
        x = self.avgpool(self.features(x)).flatten(1)
        return self.classifier(x)
