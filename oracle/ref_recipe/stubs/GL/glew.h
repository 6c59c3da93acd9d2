// TEST INFRASTRUCTURE — empty stand-in: there is no graphics stack on either box.
#pragma once
