// TEST INFRASTRUCTURE: stand-in for the un-vendored OWL header of this name (nothing of it is used by loadOBJ).
#pragma once
