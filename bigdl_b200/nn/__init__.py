from .module import (AbstractModule, Container, DynamicContainer, Echo, Graph, Identity, Input, Model, Module, Node,
                     Sequential, StaticGraph, TensorModule)
