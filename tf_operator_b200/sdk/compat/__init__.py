"""Stand-ins for third-party packages that scripts written against the reference SDK import
next to ``kubeflow.tfjob`` (today: the four ``kubernetes.client`` models a job spec is built
from). Put this directory on ``sys.path`` only when the real package is absent."""
import os

PATH = os.path.dirname(os.path.abspath(__file__))
