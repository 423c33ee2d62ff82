#pragma once
// TEST INFRASTRUCTURE ONLY -- included by the reference, not used by any code that is instantiated.
