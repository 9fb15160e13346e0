/* TEST INFRASTRUCTURE: empty stand-in for the MSVC header the reference's stdafx.h:11 includes */
