"""The Allocate / ListAndWatch / Go-semantics parity suite, collected a second time under the
`gpu` marker so that the GPU box's `pytest -m gpu` run covers SURVEY 8(a) rows A1-A12 too, not
only the kernel (K1). Nothing here needs a device: these are the same test functions the CPU
suite runs (golden cases G1-G10 / H* / F*, seeded random pod sequences, hypothesis properties,
mutated wires, the C++ reference model against the Python one), imported from their modules."""
import pytest

from tests.test_allocate_parity import *      # noqa: F401,F403
from tests.test_oracle import *               # noqa: F401,F403
from tests.test_property_parity import *      # noqa: F401,F403
from tests.test_ref_cpp import *              # noqa: F401,F403
from tests.test_table_law import *            # noqa: F401,F403
from tests.test_wire_protobuf import *        # noqa: F401,F403

pytestmark = pytest.mark.gpu
