/* TEST INFRASTRUCTURE: empty stand-in for the Windows SDK header the reference's targetver.h:8 includes */
